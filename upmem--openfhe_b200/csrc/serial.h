// serial.h — the reference's binary wire format of one DCRTPoly, host side (no CUDA in this header).
//
// What it restates (citations relative to the reference tree; cereal is the vendored third-party/cereal):
//   Serial::Serialize / Deserialize (SerType::BINARY)       src/core/include/utils/serial.h:99-115
//     = cereal::PortableBinary{Output,Input}Archive: one endianness byte (1 = little), then the object;
//       a class version (u32) precedes the FIRST object of each versioned class only
//   DCRTPolyImpl::save   "v" towers, "f" format, "p" params  lattice/hal/default/dcrtpoly.h:349-354
//   PolyImpl::save       "v" unique_ptr<values>, "f", "p"    lattice/hal/default/poly.h:322-327
//   NativeVectorT::save  size u64, N raw u64, modulus        math/hal/intnat/mubintvecnat.h:666-675
//   NativeIntegerT::save u64                                 math/hal/intnat/ubintnat.h:1833-1836
//   ILParamsImpl / ElemParams::save  co rd cm ru bm br       lattice/ilparams.h:141-144, lattice/elemparams.h:223-231
//   ILDCRTParams::save   base, "p" tower params, "m"         lattice/ildcrtparams.h:350-355
//   bigintdyn::ubint::save  limbs vector<u64>, MSB u32       math/hal/bigintdyn/ubintdyn.h:900-904 (MATHBACKEND 4)
// cereal's smart-pointer encoding (types/memory.hpp, types/polymorphic.hpp): a pointer to a polymorphic class whose
// dynamic type is its static type is written as polymorphic_id = 0x40000000; a unique_ptr then carries a `valid` byte,
// a shared_ptr a u32 id whose top bit marks the first occurrence (the pointee follows) — later occurrences are the id
// alone.  Tower params are written once, inside the tower that uses them, and referenced by id from ILDCRTParams.
//
// The tower payload (N raw little-endian u64) sits at byte offsets that are not 8-aligned (34 for tower 0), so the
// device side moves it with byte-granular 2-D copies straight between HBM and the blob (runtime.cu).
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace fhe {
namespace serial {

constexpr uint32_t POLY_ID_SAME_TYPE = 0x40000000u;   // cereal detail::msb2_32bit
constexpr uint32_t PTR_FIRST         = 0x80000000u;   // cereal detail::msb_32bit
constexpr uint32_t CLASS_VERSION     = 1;             // every SerializedVersion() on this path returns 1

// ---- writer ---------------------------------------------------------------------------------------------------
struct Out {
    std::vector<uint8_t> b;
    template <class T> void put(T v) {
        uint8_t tmp[sizeof(T)];
        std::memcpy(tmp, &v, sizeof(T));
        b.insert(b.end(), tmp, tmp + sizeof(T));
    }
};

// little-endian limbs of prod(q); {0} for an empty product is never needed (T >= 1)
inline std::vector<uint64_t> big_product(const std::vector<uint64_t>& q) {
    std::vector<uint64_t> v{1};
    for (uint64_t m : q) {
        unsigned __int128 carry = 0;
        for (auto& limb : v) {
            unsigned __int128 t = (unsigned __int128)limb * m + carry;
            limb  = (uint64_t)t;
            carry = t >> 64;
        }
        if (carry) v.push_back((uint64_t)carry);
    }
    return v;
}
inline uint32_t big_msb(const std::vector<uint64_t>& v) {
    for (size_t i = v.size(); i-- > 0;)
        if (v[i]) return (uint32_t)(64 * i + 64 - __builtin_clzll(v[i]));
    return 0;
}
inline void put_big(Out& o, const std::vector<uint64_t>& limbs, bool& seen_big) {
    if (!seen_big) { o.put<uint32_t>(CLASS_VERSION); seen_big = true; }
    o.put<uint64_t>(limbs.size());
    for (uint64_t l : limbs) o.put<uint64_t>(l);
    o.put<uint32_t>(big_msb(limbs));
}

// The blob of one poly as T+1 header segments around T payload holes.
struct Layout {
    size_t bytes = 0;                        // whole blob
    std::vector<size_t> data_off;            // [T] byte offset of tower t's N*8 payload
    std::vector<size_t> seg_off;             // [T+1] where each header segment starts
    std::vector<std::vector<uint8_t>> seg;   // [T+1] header bytes; seg[t] precedes tower t's payload, seg[T] ends the blob
    std::vector<std::pair<uint32_t, uint32_t>> format_at;   // (segment, offset in it) of every Format field
};

// Canonical stream of DCRTPoly(params, format) whose towers share their params with ILDCRTParams — what every
// constructor of the reference produces (dcrtpoly-impl.h:60-75).  Tower params carry no big modulus / root
// (ILDCRTParams(order, moduli, roots), ildcrtparams.h:140-150).
inline Layout make_layout(uint32_t N, const std::vector<uint64_t>& q, const std::vector<uint64_t>& psi) {
    const uint32_t T = (uint32_t)q.size();
    Layout L;
    L.seg.resize(T + 1);
    L.seg_off.resize(T + 1);
    L.data_off.resize(T);
    bool seen_big = false;
    size_t pos = 0;
    Out o;
    o.put<uint8_t>(1);                       // PortableBinary: little endian
    o.put<uint32_t>(CLASS_VERSION);          // DCRTPolyImpl
    o.put<uint64_t>(T);                      // "v": vector<PolyType>
    for (uint32_t t = 0; t < T; t++) {
        if (t == 0) o.put<uint32_t>(CLASS_VERSION);           // PolyImpl
        o.put<uint32_t>(POLY_ID_SAME_TYPE);                   // "v": unique_ptr<NativeVector>
        o.put<uint8_t>(1);                                    //   valid
        if (t == 0) o.put<uint32_t>(CLASS_VERSION);           // NativeVector
        o.put<uint64_t>(N);
        L.seg_off[t] = pos;
        pos += o.b.size();
        L.data_off[t] = pos;
        L.seg[t].swap(o.b);
        pos += (size_t)N * 8;
        o.b.clear();
        if (t == 0) o.put<uint32_t>(CLASS_VERSION);           // NativeInteger (the vector's modulus is the first one)
        o.put<uint64_t>(q[t]);
        L.format_at.push_back({t + 1, (uint32_t)o.b.size()});
        o.put<uint32_t>(0);                                   // "f" of the tower
        o.put<uint32_t>(POLY_ID_SAME_TYPE);                   // "p": shared_ptr<ILNativeParams>
        o.put<uint32_t>(PTR_FIRST | (t + 1));
        if (t == 0) { o.put<uint32_t>(CLASS_VERSION); o.put<uint32_t>(CLASS_VERSION); }   // ILParamsImpl, ElemParams<NativeInteger>
        o.put<uint32_t>(2 * N);                               // co
        o.put<uint32_t>(N);                                   // rd
        o.put<uint64_t>(q[t]);                                // cm
        o.put<uint64_t>(psi.empty() ? 0 : psi[t]);            // ru
        o.put<uint64_t>(0);                                   // bm
        o.put<uint64_t>(0);                                   // br
    }
    L.format_at.push_back({T, (uint32_t)o.b.size()});
    o.put<uint32_t>(0);                                       // "f" of the DCRTPoly
    o.put<uint32_t>(POLY_ID_SAME_TYPE);                       // "p": shared_ptr<ILDCRTParams<BigInteger>>
    o.put<uint32_t>(PTR_FIRST | (T + 1));
    o.put<uint32_t>(CLASS_VERSION);                           // ILDCRTParams
    o.put<uint32_t>(CLASS_VERSION);                           // ElemParams<BigInteger>
    o.put<uint32_t>(2 * N);
    o.put<uint32_t>(N);
    put_big(o, big_product(q), seen_big);                     // cm = prod q_i (ildcrtparams.h RecalculateModulus)
    put_big(o, {0}, seen_big);                                // ru
    put_big(o, {0}, seen_big);                                // bm
    put_big(o, {0}, seen_big);                                // br
    o.put<uint64_t>(T);                                       // "p": vector<shared_ptr<ILNativeParams>>
    for (uint32_t t = 0; t < T; t++) {
        o.put<uint32_t>(POLY_ID_SAME_TYPE);
        o.put<uint32_t>(t + 1);                               // already written: id only
    }
    put_big(o, {0}, seen_big);                                // "m": original modulus
    L.seg_off[T] = pos;
    pos += o.b.size();
    L.seg[T].swap(o.b);
    L.bytes = pos;
    // the loop above recorded the tower-t format inside the segment that FOLLOWS tower t's payload (t + 1), and the
    // poly's own format in the last segment
    return L;
}

inline void write_headers(const Layout& L, uint8_t* blob, uint32_t format) {
    for (size_t s = 0; s < L.seg.size(); s++) std::memcpy(blob + L.seg_off[s], L.seg[s].data(), L.seg[s].size());
    for (auto& f : L.format_at) std::memcpy(blob + L.seg_off[f.first] + f.second, &format, 4);
}

// ---- reader ---------------------------------------------------------------------------------------------------
// Accepts any stream the reference's save functions can produce for a DCRTPoly over native towers (shared or
// separately owned tower params, class versions <= 1) and checks it against the ring it is loaded into.
struct Parsed {
    uint32_t format = 0;
    std::vector<size_t> data_off;   // [T]
    size_t consumed = 0;
};

struct In {
    const uint8_t* p;
    size_t n, pos = 0;
    std::string err;
    bool ok() const { return err.empty(); }
    template <class T> T get() {
        T v{};
        if (!ok()) return v;
        if (n - pos < sizeof(T)) { err = "stream ends inside a field"; return v; }
        std::memcpy(&v, p + pos, sizeof(T));
        pos += sizeof(T);
        return v;
    }
    void skip(size_t k) {
        if (!ok()) return;
        if (n - pos < k) { err = "stream ends inside the tower payload"; return; }
        pos += k;
    }
    void version(bool& seen, const char* cls) {
        if (seen || !ok()) return;
        seen = true;
        uint32_t v = get<uint32_t>();
        if (ok() && v > CLASS_VERSION) err = std::string("serialized ") + cls + " is from a later version of the library";
    }
    void poly_id(const char* what) {
        uint32_t id = get<uint32_t>();
        if (ok() && id != POLY_ID_SAME_TYPE) err = std::string(what) + ": null or derived-type pointer in the stream";
    }
};

struct NativeParams { uint32_t co = 0, rd = 0; uint64_t cm = 0, ru = 0, bm = 0, br = 0; };
struct Seen { bool poly = false, vec = false, nint = false, ilp = false, epn = false, big = false; };

inline std::vector<uint64_t> get_big(In& in, Seen& s) {
    in.version(s.big, "BigInteger");
    uint64_t limbs = in.get<uint64_t>();
    std::vector<uint64_t> v;
    if (in.ok() && limbs > (in.n - in.pos) / 8) in.err = "BigInteger limb count exceeds the stream";
    for (uint64_t i = 0; in.ok() && i < limbs; i++) v.push_back(in.get<uint64_t>());
    in.get<uint32_t>();   // MSB, recomputed by the reader
    return v;
}
inline uint64_t get_nint(In& in, Seen& s) {
    in.version(s.nint, "NativeInteger");
    return in.get<uint64_t>();
}
// shared_ptr<ILNativeParams>: returns the pointee (read here or looked up by id)
inline NativeParams get_native_params(In& in, Seen& s, std::map<uint32_t, NativeParams>& ids) {
    NativeParams np;
    in.poly_id("tower params");
    uint32_t id = in.get<uint32_t>();
    if (!in.ok()) return np;
    if (!(id & PTR_FIRST)) {
        auto it = ids.find(id);
        if (it == ids.end()) { in.err = "tower params refer to an id the stream has not defined"; return np; }
        return it->second;
    }
    in.version(s.ilp, "ILParamsImpl");
    in.version(s.epn, "ElemParams");
    np.co = in.get<uint32_t>();
    np.rd = in.get<uint32_t>();
    np.cm = get_nint(in, s);
    np.ru = get_nint(in, s);
    np.bm = get_nint(in, s);
    np.br = get_nint(in, s);
    ids[id & ~PTR_FIRST] = np;
    return np;
}

inline bool check_params(In& in, const NativeParams& np, uint32_t N, uint64_t q, const uint64_t* psi, uint32_t t) {
    if (!in.ok()) return false;
    if (np.co != 2 * N || np.rd != N) { in.err = "tower " + std::to_string(t) + ": ring dimension differs from the ring's"; return false; }
    if (np.cm != q) { in.err = "tower " + std::to_string(t) + ": modulus differs from the ring's"; return false; }
    if (psi && np.ru != *psi) { in.err = "tower " + std::to_string(t) + ": root of unity differs from the ring's"; return false; }
    return true;
}

// One stream = one DCRTPoly.  psi may be empty (ring without roots: the roots in the stream are not checked).
inline bool parse(const uint8_t* blob, size_t len, uint32_t N, const std::vector<uint64_t>& q,
                  const std::vector<uint64_t>& psi, Parsed& out, std::string& err) {
    const uint32_t T = (uint32_t)q.size();
    In in{blob, len};
    Seen s;
    std::map<uint32_t, NativeParams> ids;
    auto bad = [&](const std::string& m) { err = in.ok() ? m : in.err; return false; };
    if (in.get<uint8_t>() != 1) return bad("not a little-endian portable binary stream");
    bool seen_dcrt = false;
    in.version(seen_dcrt, "DCRTPolyImpl");
    uint64_t towers = in.get<uint64_t>();
    if (!in.ok() || towers != T) return bad("tower count differs from the ring's");
    out.data_off.assign(T, 0);
    std::vector<uint32_t> tf(T);
    for (uint32_t t = 0; t < T; t++) {
        in.version(s.poly, "PolyImpl");
        in.poly_id("tower values");
        uint8_t valid = in.get<uint8_t>();
        if (in.ok() && valid != 1) return bad("tower " + std::to_string(t) + " has no values");
        in.version(s.vec, "NativeVector");
        uint64_t size = in.get<uint64_t>();
        if (!in.ok() || size != N) return bad("tower " + std::to_string(t) + ": length differs from the ring dimension");
        out.data_off[t] = in.pos;
        in.skip((size_t)N * 8);
        uint64_t mod = get_nint(in, s);
        if (!in.ok() || mod != q[t]) return bad("tower " + std::to_string(t) + ": vector modulus differs from the ring's");
        tf[t] = in.get<uint32_t>();
        NativeParams np = get_native_params(in, s, ids);
        if (!check_params(in, np, N, q[t], psi.empty() ? nullptr : &psi[t], t)) return bad("");
    }
    out.format = in.get<uint32_t>();
    if (!in.ok() || out.format > 1) return bad("format field is neither EVALUATION nor COEFFICIENT");
    for (uint32_t t = 0; t < T; t++)
        if (tf[t] != out.format) return bad("tower " + std::to_string(t) + ": format differs from the poly's");
    in.poly_id("ILDCRTParams");
    uint32_t id = in.get<uint32_t>();
    if (!in.ok()) return bad("");
    if (!(id & PTR_FIRST)) return bad("ILDCRTParams refer to an id the stream has not defined");
    bool seen_ildcrt = false, seen_epb = false;
    in.version(seen_ildcrt, "ILDCRTParams");
    in.version(seen_epb, "ElemParams");
    uint32_t co = in.get<uint32_t>(), rd = in.get<uint32_t>();
    if (!in.ok() || co != 2 * N || rd != N) return bad("ILDCRTParams: ring dimension differs from the ring's");
    std::vector<uint64_t> cm = get_big(in, s);
    while (cm.size() > 1 && cm.back() == 0) cm.pop_back();
    if (in.ok() && cm != big_product(q)) return bad("ILDCRTParams: composite modulus is not the product of the ring's moduli");
    get_big(in, s);   // ru
    get_big(in, s);   // bm
    get_big(in, s);   // br
    uint64_t cnt = in.get<uint64_t>();
    if (!in.ok() || cnt != T) return bad("ILDCRTParams: tower count differs from the ring's");
    for (uint32_t t = 0; t < T; t++) {
        NativeParams np = get_native_params(in, s, ids);
        if (!check_params(in, np, N, q[t], psi.empty() ? nullptr : &psi[t], t)) return bad("");
    }
    get_big(in, s);   // original modulus
    if (!in.ok()) return bad("");
    out.consumed = in.pos;
    return true;
}

}  // namespace serial
}  // namespace fhe
