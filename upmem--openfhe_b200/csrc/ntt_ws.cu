// ntt_ws.cu — warp-specialised persistent passes of the batched negacyclic NTT / INTT (N = 2^14, 2^15, 2^16).
//
// Same arithmetic, same two-pass decomposition N = N1 x 256 and same register steps (ntt_steps.cuh) as ntt.cu, so the
// outputs are the reference's (transformnat-impl.h:300-354, 492-552) bit for bit.  What changes is how tiles reach the
// butterfly warps.  In ntt.cu a CTA is one tile: its warps issue their own global loads and wait for them, and a
// quarter of all warp time is that wait (ncu, round 1).  Here one persistent CTA per SM holds
//   * one PRODUCER warp that walks the CTA's range of tiles and moves each tile from HBM into a ring of K shared-memory
//     slots with bulk asynchronous copies (cp.async.bulk ... mbarrier::complete_tx — the TMA engine does the transfer,
//     no register, no LSU instruction and no waiting warp is spent on it), and
//   * G COMPUTE groups of four warps; a group takes every G-th tile of the ring: waits for the slot's `full` barrier,
//     runs the first register step out of shared memory, exchanges in place, runs the second step, and
//       - strided pass: stores its rows straight from registers (64..256-byte row segments), then releases the slot;
//       - contig pass : writes the finished tile back into the slot and one thread hands it to the TMA engine
//                       (cp.async.bulk shared -> global, 2 KB per sub-NTT), releasing the slot when that has been read.
// So global loads are always K - G tiles ahead of the butterflies, every resident warp but one is a butterfly warp that
// finds its operands in shared memory, and a tile costs two group barriers.  Tiles are walked in memory order (strided
// pass) / batch-fastest order (contig pass: consecutive tiles of a CTA share their 255 twiddles, which then come from L1).
//
// Tile = 2048 coefficients (16 KB) in both passes:
//   strided: R = N/256 rows x C = 2048/R columns of one (batch item, tower);
//   contig : the SAME 256-point block of 8 different batch items (twiddle loads shared), each padded by 16 B in the slot
//            so that the per-thread 128-byte runs of the second step are bank-conflict free with 16-byte accesses.
#include "ntt_steps.cuh"
#include "kernels.h"

#include <cstdlib>

namespace fhe {
namespace ws {

constexpr int GT = 128;                    // threads per compute group
constexpr int SLOT_WORDS = 2176;           // 2048 coefficients + padding (strided C = 8: 128 words; contig: 8 x 2)
constexpr int SLOT_BYTES = SLOT_WORDS * 8;
constexpr int CROW = 258;                  // words between the sub-NTTs of a contig tile (256 + 16 B)

// G compute groups per CTA and a ring of K = G + 3 slots (the producer runs up to 3 tiles ahead of the slowest group).
// G = 6: 25 warps, 72 registers per thread; G = 5: 21 warps, 80 registers (FHE_B200_NTT_WS_GROUPS picks, default 6).
template <int G_>
struct Cfg {
    static constexpr int G = G_, K = G_ + 3;
    static constexpr int THREADS = G * GT + 32;       // + the producer warp
    static constexpr int BAR_OFF = K * SLOT_BYTES;    // full[K], empty[K]
    static constexpr int KQ_OFF = BAR_OFF + 2 * K * 8;
    static constexpr int SMEM = KQ_OFF + G * 16 * 8;  // + per-group {0, q, .., 15q}
};

__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(u32 bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 bar, u32 parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "LAB_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
        "@P1 bra DONE;\n\t"
        "bra LAB_WAIT;\n\t"
        "DONE:\n\t"
        "}" ::"r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive(u32 bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(u32 bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// HBM -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_g2s(u32 dst, const void* src, u32 bytes, u32 bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// shared -> HBM (bulk group of the issuing thread)
__device__ __forceinline__ void bulk_s2g(void* dst, u32 src, u32 bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// generic-proxy writes to shared memory become visible to the asynchronous proxy (the TMA engine)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void group_bar(u32 g) { asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(GT) : "memory"); }

struct Ring {
    u64* slots;
    u32 full0, empty0;          // shared addresses of full[0], empty[0]
    __device__ __forceinline__ u32 full(u32 s) const { return full0 + 8 * s; }
    __device__ __forceinline__ u32 empty(u32 s) const { return empty0 + 8 * s; }
};

// common prologue: barriers, this CTA's contiguous range of tiles
template <class CF>
__device__ __forceinline__ Ring ring_setup(unsigned char* sm, u32 empty_count, u32 tiles, u32& t0, u32& t1) {
    constexpr int K = CF::K;
    Ring r;
    r.slots = reinterpret_cast<u64*>(sm);
    r.full0 = smem_u32(sm + CF::BAR_OFF);
    r.empty0 = r.full0 + 8 * K;
    if (threadIdx.x == 0) {
        for (u32 s = 0; s < (u32)K; s++) { mbar_init(r.full(s), 1); mbar_init(r.empty(s), empty_count); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const u32 per = (tiles + gridDim.x - 1) / gridDim.x;
    t0 = blockIdx.x * per;
    t1 = t0 + per < tiles ? t0 + per : tiles;
    if (t0 > t1) t0 = t1;
    return r;
}

// ------------------------------------------------------------------------------------------
// Strided pass: R = 2^RL point sub-NTTs over elements 256 apart; tile = R rows x C columns.
// ------------------------------------------------------------------------------------------
template <int RL>
struct SCfg {
    static constexpr int A = 4, B = RL - 4, R = 1 << RL, C = 2048 / R;       // RL = 6, 7, 8 -> C = 32, 16, 8
    static constexpr int PAD = C < 16 ? C : 0;        // one extra row per 2^B rows keeps both access patterns conflict free
    static constexpr int GB = 16 >> B;
    static constexpr int NCG = 256 / C;               // tiles per (batch item, tower)
    __device__ static __forceinline__ u32 sidx(u32 row, u32 c) { return row * C + c + (row >> B) * PAD; }
    static_assert(RL >= 6 && RL <= 8, "ring size not covered");
    static_assert(R * C + (R >> B) * PAD <= SLOT_WORDS, "slot too small");
};

template <int RL, bool INV, int G>
__global__ void __launch_bounds__(Cfg<G>::THREADS, 1)
ntt_strided_ws_kernel(u64* data, const u64* in, const ulonglong2* __restrict__ twAll,
                      const TowerMod* __restrict__ mods, const ulonglong2* __restrict__ invc,
                      u32 T, u32 logN, u32 outT, u32 inT, u32 tiles) {
    using S = SCfg<RL>;
    using CF = Cfg<G>;
    constexpr int B = S::B, C = S::C, K = CF::K;
    extern __shared__ __align__(128) unsigned char smraw[];
    u32 t0, t1;
    const Ring ring = ring_setup<CF>(smraw, 4, tiles, t0, t1);
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == G * 4) {
        // ---- producer: one bulk copy per row of the tile, rows spread over the lanes ----
        for (u32 t = t0, i = 0; t < t1; t++, i++) {
            const u32 s = i % K;
            if (i >= (u32)K) mbar_wait(ring.empty(s), ((i / K) & 1) ^ 1);
            const u32 cg = t % S::NCG, u = t / S::NCG, tower = u % T, b = u / T;
            const u64* src = in + (((size_t)b * inT + tower) << logN) + cg * C;
            const u32 dst = smem_u32(ring.slots + (size_t)s * SLOT_WORDS);
            if (lane == 0) mbar_arrive_expect_tx(ring.full(s), 2048 * 8);
            __syncwarp();
#pragma unroll 4
            for (u32 row = lane; row < (u32)S::R; row += 32)
                bulk_g2s(dst + S::sidx(row, 0) * 8, src + ((size_t)row << 8), C * 8, ring.full(s));
        }
        return;
    }

    const u32 g = warp >> 2, tid = threadIdx.x & (GT - 1);
    for (u32 i = g; t0 + i < t1; i += G) {
        const u32 t = t0 + i, s = i % K;
        const u32 cg = t % S::NCG, u = t / S::NCG, tower = u % T, b = u / T;
        const u64 q = mods[tower].q, nq = 0ull - q;
        const QMul Q = make_qmul(q);
        const ulonglong2* tw = twAll + ((size_t)tower << logN);
        u64* base = data + (((size_t)b * outT + tower) << logN) + cg * C;
        u64* sm = ring.slots + (size_t)s * SLOT_WORDS;
        mbar_wait(ring.full(s), (i / K) & 1);
        if (!INV) {
            {   // step A: the top four bits of the row index live in registers; exchange in place
                const u32 c = tid % C, ulow = tid / C;
                u64 x[16];
#pragma unroll
                for (int e = 0; e < 16; e++) x[e] = sm[S::sidx((e << B) + ulow, c)];
                fwd_step<4, 1>(x, tw, 1u, Q, nq);
#pragma unroll
                for (int e = 0; e < 16; e++) sm[S::sidx((e << B) + ulow, c)] = x[e];
            }
            group_bar(g);
#pragma unroll
            for (int gi = 0; gi < S::GB; gi++) {
                const u32 item = tid + gi * GT, c = item % C, uhi = item / C;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = sm[S::sidx((uhi << B) + e, c)];
                fwd_step<B, FwdStep<4, 1>::OUT>(x, tw, 16u + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) base[((size_t)((uhi << B) + e) << 8) + c] = x[e];
            }
        } else {
            const ulonglong2 sc = invc[2 * tower], w1 = invc[2 * tower + 1];
#pragma unroll
            for (int gi = 0; gi < S::GB; gi++) {
                const u32 item = tid + gi * GT, c = item % C, uhi = item / C;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = sm[S::sidx((uhi << B) + e, c)];
                inv_step<B, INV_NORM, INV_NORM>(x, tw, 16u + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) sm[S::sidx((uhi << B) + e, c)] = x[e];
            }
            group_bar(g);
            {
                const u32 c = tid % C, ulow = tid / C;
                u64 x[16];
#pragma unroll
                for (int e = 0; e < 16; e++) x[e] = sm[S::sidx((e << B) + ulow, c)];
                inv_step_last<4, INV_NORM>(x, tw, sc, w1, Q, nq);
#pragma unroll
                for (int e = 0; e < 16; e++) base[((size_t)((e << B) + ulow) << 8) + c] = x[e];
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(ring.empty(s));      // this warp's reads of the slot are done
    }
}

// ------------------------------------------------------------------------------------------
// Contig pass: 256-point sub-NTTs over contiguous blocks (transform stages L-8..L-1).
// Tile = block h of tower `tower` of 8 consecutive batch items.
// ------------------------------------------------------------------------------------------
template <bool INV, int G>
__global__ void __launch_bounds__(Cfg<G>::THREADS, 1)
ntt_contig_ws_kernel(u64* data, const u64* in, const ulonglong2* __restrict__ twAll,
                     const TowerMod* __restrict__ mods, u32 T, u32 logN, u32 batch, u32 outT, u32 inT, u32 tiles) {
    using CF = Cfg<G>;
    constexpr int K = CF::K;
    extern __shared__ __align__(128) unsigned char smraw[];
    u32 t0, t1;
    const Ring ring = ring_setup<CF>(smraw, 1, tiles, t0, t1);
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const u32 N1 = 1u << (logN - 8), nbg = (batch + 7) >> 3;

    if (warp == G * 4) {
        for (u32 t = t0, i = 0; t < t1; t++, i++) {
            const u32 s = i % K;
            if (i >= (u32)K) mbar_wait(ring.empty(s), ((i / K) & 1) ^ 1);
            const u32 bg = t % nbg, u = t / nbg, h = u & (N1 - 1), tower = u >> (logN - 8);
            const u32 nv = batch - bg * 8 < 8 ? batch - bg * 8 : 8;
            const u32 dst = smem_u32(ring.slots + (size_t)s * SLOT_WORDS);
            if (lane == 0) mbar_arrive_expect_tx(ring.full(s), nv * 2048);
            __syncwarp();
            if (lane < nv) {
                const u64* src = in + (((size_t)(bg * 8 + lane) * inT + tower) << logN) + ((size_t)h << 8);
                bulk_g2s(dst + lane * (CROW * 8), src, 2048, ring.full(s));
            }
        }
        return;
    }

    const u32 g = warp >> 2, tid = threadIdx.x & (GT - 1);
    u64* kq = reinterpret_cast<u64*>(smraw + CF::KQ_OFF) + g * 16;      // {0, q, .., 15q} for the closing range reduction
    for (u32 i = g; t0 + i < t1; i += G) {
        const u32 t = t0 + i, s = i % K;
        const u32 bg = t % nbg, u = t / nbg, h = u & (N1 - 1), tower = u >> (logN - 8);
        const u32 nv = batch - bg * 8 < 8 ? batch - bg * 8 : 8;
        const TowerMod tm = mods[tower];
        const u64 q = tm.q, nq = 0ull - q;
        const QMul Q = make_qmul(q);
        const ulonglong2* tw = twAll + ((size_t)tower << logN);
        u64* sm = ring.slots + (size_t)s * SLOT_WORDS;
        mbar_wait(ring.full(s), (i / K) & 1);
        if (!INV && tid < 16) kq[tid] = tid * q;
        const u32 ca = tid >> 4, ulow = tid & 15;       // first-step mapping of the forward transform: lanes along u
        const u32 cb = tid & 7, uhi = tid >> 3;         // second-step mapping: 16 consecutive u per thread
        if (nv < 8) {                                    // last batch group: absent items compute on zeros
            const u32 cz = INV ? cb : ca;
            if (cz >= nv) {
#pragma unroll
                for (int e = 0; e < 16; e++) sm[cz * CROW + (INV ? uhi * 16 + e : e * 16 + ulow)] = 0ull;
            }
        }
        if (!INV) {
            {
                u64 x[16];
#pragma unroll
                for (int e = 0; e < 16; e++) x[e] = sm[ca * CROW + e * 16 + ulow];
                fwd_step<4, MAXB>(x, tw, N1 + h, Q, nq);          // entry: whatever the strided pass left, < 16q
#pragma unroll
                for (int e = 0; e < 16; e++) sm[ca * CROW + e * 16 + ulow] = x[e];
            }
            group_bar(g);
            {
                u64 x[16];
                ulonglong2* row = reinterpret_cast<ulonglong2*>(sm + cb * CROW + uhi * 16);
#pragma unroll
                for (int k = 0; k < 8; k++) { ulonglong2 v = row[k]; x[2 * k] = v.x; x[2 * k + 1] = v.y; }
                fwd_step<4, FwdStep<4, MAXB>::OUT>(x, tw, ((N1 + h) << 4) + uhi, Q, nq);
#pragma unroll
                for (int k = 0; k < 8; k++)
                    row[k] = make_ulonglong2(reduce16(x[2 * k], kq, tm.rsh, tm.rinv, q), reduce16(x[2 * k + 1], kq, tm.rsh, tm.rinv, q));
            }
        } else {
            {
                u64 x[16];
                ulonglong2* row = reinterpret_cast<ulonglong2*>(sm + cb * CROW + uhi * 16);
#pragma unroll
                for (int k = 0; k < 8; k++) { ulonglong2 v = row[k]; x[2 * k] = v.x; x[2 * k + 1] = v.y; }
                inv_step<4, 1, INV_NORM>(x, tw, ((N1 + h) << 4) + uhi, Q, nq);
#pragma unroll
                for (int k = 0; k < 8; k++) row[k] = make_ulonglong2(x[2 * k], x[2 * k + 1]);
            }
            group_bar(g);
            {
                u64 x[16];
#pragma unroll
                for (int e = 0; e < 16; e++) x[e] = sm[ca * CROW + e * 16 + ulow];
                inv_step<4, INV_NORM, INV_NORM>(x, tw, N1 + h, Q, nq);
#pragma unroll
                for (int e = 0; e < 16; e++) sm[ca * CROW + e * 16 + ulow] = x[e];   // lazily reduced, [0, INV_NORM*q)
            }
        }
        fence_async_smem();
        group_bar(g);
        if (tid == 0) {      // the finished tile goes back to HBM through the TMA engine; the slot is free once it has been read
            const u32 src = smem_u32(sm);
            for (u32 c = 0; c < nv; c++)
                bulk_s2g(data + (((size_t)(bg * 8 + c) * outT + tower) << logN) + ((size_t)h << 8), src + c * (CROW * 8), 2048);
            bulk_commit();
            bulk_wait_read();
            mbar_arrive(ring.empty(s));
        }
    }
}

// ------------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------------
static int sm_count() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

static int groups() {                      // compute groups per CTA: 6 (default) or 5
    static int g = 0;
    if (!g) {
        const char* e = getenv("FHE_B200_NTT_WS_GROUPS");
        g = e && atoi(e) == 5 ? 5 : 6;
    }
    return g;
}

template <int RL, bool INV, int G>
static cudaError_t launch_strided(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    using CF = Cfg<G>;
    const size_t tiles = batch * r.T * SCfg<RL>::NCG;
    if (tiles > 0xffffffffull) return cudaErrorInvalidValue;
    auto kern = ntt_strided_ws_kernel<RL, INV, G>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SMEM);
    if (e != cudaSuccess) return e;
    unsigned grid = (unsigned)std::min<size_t>((size_t)sm_count(), (tiles + G - 1) / G);
    kern<<<grid, CF::THREADS, CF::SMEM, st>>>(data, in, INV ? r.twI : r.twF, r.mods, r.invc, r.T, r.logN, outT, inT, (u32)tiles);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

template <bool INV, int G>
static cudaError_t launch_contig(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    using CF = Cfg<G>;
    const size_t tiles = ((batch + 7) / 8) * r.T * (r.N >> 8);
    if (tiles > 0xffffffffull || batch > 0xffffffffull) return cudaErrorInvalidValue;
    auto kern = ntt_contig_ws_kernel<INV, G>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SMEM);
    if (e != cudaSuccess) return e;
    unsigned grid = (unsigned)std::min<size_t>((size_t)sm_count(), (tiles + G - 1) / G);
    kern<<<grid, CF::THREADS, CF::SMEM, st>>>(data, in, INV ? r.twI : r.twF, r.mods, r.T, r.logN, (u32)batch, outT, inT, (u32)tiles);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

template <bool INV, int G>
static cudaError_t launch_strided_any(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    switch (r.logN) {
        case 14: return launch_strided<6, INV, G>(r, data, in, outT, inT, batch, st);
        case 15: return launch_strided<7, INV, G>(r, data, in, outT, inT, batch, st);
        case 16: return launch_strided<8, INV, G>(r, data, in, outT, inT, batch, st);
    }
    return cudaErrorInvalidValue;
}

template <int G>
static cudaError_t run(const RingDev& r, u64* d, const u64* s, size_t batch, bool inverse, u32 outT, u32 inT, cudaStream_t st) {
    cudaError_t e;
    if (!inverse) {
        e = launch_strided_any<false, G>(r, d, s, outT, inT, batch, st);
        if (e != cudaSuccess) return e;
        return launch_contig<false, G>(r, d, d, outT, outT, batch, st);
    }
    e = launch_contig<true, G>(r, d, s, outT, inT, batch, st);
    if (e != cudaSuccess) return e;
    return launch_strided_any<true, G>(r, d, d, outT, outT, batch, st);
}

}  // namespace ws

// FHE_B200_NTT_WS = 0: never; 1: every transform of a ring the kernels cover (tests); unset: when there are enough tiles to
// keep the ring of every SM busy.
bool ntt_ws_applies(const RingDev& r, size_t batch) {
    if (r.logN < 14 || r.logN > 16) return false;
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("FHE_B200_NTT_WS");
        mode = e && *e ? (atoi(e) == 2 ? 2 : atoi(e) ? 1 : 0) : 0;        // default off: see DESIGN.md §9 (round 2)
    }
    if (mode != 2) return mode == 1;
    return batch * r.T * (r.N >> 11) >= (size_t)ws::sm_count() * 6 * 4;           // 2048-coefficient tiles
}

cudaError_t launch_ntt_ws(const RingDev& r, uint64_t* data, size_t batch, bool inverse, cudaStream_t st,
                          const uint64_t* in, uint32_t outT, uint32_t inT) {
    u64* d = reinterpret_cast<u64*>(data);
    const u64* s = reinterpret_cast<const u64*>(in);
    return ws::groups() == 5 ? ws::run<5>(r, d, s, batch, inverse, outT, inT, st) : ws::run<6>(r, d, s, batch, inverse, outT, inT, st);
}

}  // namespace fhe
