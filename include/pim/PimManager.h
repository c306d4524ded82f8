/*
 * pim/PimManager.h — the runtime object of the intercept surface, over the B200 C ABI.
 *
 * Same class name, header path and public methods as the reference's UPMEM host runtime
 * (src/core/include/pim/PimManager.h:21-127, src/core/pim/host/PimManager.cpp:5-96), so the
 * SetPim/GetPim plumbing in OpenFHE's Ciphertext / DCRTPoly / Poly / NativeVector compiles
 * untouched.  What changed behind it:
 *
 *   reference (UPMEM)                                 here (B200)
 *   dpu_alloc(nr_dpus, profile) + load BOOT           fhe_init(device, opts): stream pool + HBM arena
 *   one DpuMemory (64 MB MRAM) per DPU                one device heap (default 1 GiB, `heap_mb=`) + DpuMemory
 *   copy_to_pim type 0: scatter size/nr_dpus words    one cudaMemcpyAsync H2D of `size` words at `offset`
 *   copy_to_pim type 1 to WRAM symbol "meta"          descriptor kept on the host until start_kernel
 *   load_kernel(path of a DPU binary)                 kernel id lookup (pim/kernel.h names), nothing loaded
 *   start_kernel(): dpu_launch(DPU_SYNCHRONOUS)       fhe_launch(kernel id, descriptor, heap): blocking
 *   DPU_ASSERT -> abort()                             std::runtime_error carrying fhe_last_error()
 *
 * `nr_dpus` sizes the stream pool (clamped to [1,64]); getNumDpus() returns 1 per GPU because the
 * heap is not split: descriptors built with DIV(bytes, getNumDpus()) (pim/PimData.h:281-290 in the
 * reference) therefore carry whole-vector sizes.  `profile` is the backend option string
 * ("device=0,heap_mb=1024,streams=4,arena_mb=256").  There is no CPU fallback: without a B200 the
 * constructor throws.  Header only; link with libfhe_b200.so.
 */
#ifndef _PIM_MANAGER_
#define _PIM_MANAGER_

#include "../fhe_b200.h"
#include "DpuMemory.h"
#include "common.h"
#include "kernel.h"

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#ifndef DPU_MRAM_HEAP_POINTER_NAME
#define DPU_MRAM_HEAP_POINTER_NAME "__sys_used_mram_end"
#endif

class PimManager {
public:
  static PimManager *getPim(uint32_t nr_dpus, const std::string &profile = "") {
    std::lock_guard<std::mutex> lock(mutex_);
    if (pim == nullptr) pim = new PimManager(nr_dpus, profile);
    return pim;
  }
  /* not in the reference: tears the singleton down (tests, orderly shutdown) */
  static void releasePim() {
    std::lock_guard<std::mutex> lock(mutex_);
    delete pim;
    pim = nullptr;
  }

  /* type 0: `size` 64-bit words from buf to the heap at byte `offset`;
   * type 1: `size` BYTES broadcast to a named symbol — only "meta" exists (the op descriptor). */
  void copy_to_pim(void *buf, uint32_t size, uint32_t offset, uint8_t type = 0,
                   const std::string &memory = DPU_MRAM_HEAP_POINTER_NAME) {
    switch (type) {
    case 0:
      bounds(offset, (size_t)size * 8, "copy_to_pim");
      check(fhe_h2d(ctx_, heap_ + offset, buf, (size_t)size * 8, 0), "copy_to_pim");
      check(fhe_sync(ctx_, 0), "copy_to_pim");
      break;
    case 1:
      if (memory != "meta") throw std::runtime_error("copy_to_pim: unknown device symbol '" + memory + "'");
      meta_.assign((const char *)buf, (const char *)buf + size);
      break;
    default:
      std::cout << "Not Available copy option" << std::endl;   /* PimManager.cpp:33-35 */
      break;
    }
  }

  uint32_t copy_from_pim(uint64_t *buf, uint32_t size, uint32_t offset) {
    bounds(offset, (size_t)size * 8, "copy_from_pim");
    check(fhe_d2h(ctx_, buf, heap_ + offset, (size_t)size * 8, 0), "copy_from_pim");
    check(fhe_sync(ctx_, 0), "copy_from_pim");
    return 0;
  }

  void load_kernel(const std::string &bin) {
    static const std::map<std::string, int> ids = {{BOOT, FHE_KERNEL_BOOT},
                                                   {ELEM_MOD_ADD, FHE_KERNEL_MOD_ADD},
                                                   {ELEM_MOD_MULT, FHE_KERNEL_MOD_MULT},
                                                   {ELEM_MOD_SUB, FHE_KERNEL_MOD_SUB},
                                                   {ELEM_MOD_OPS, FHE_KERNEL_MOD_OPS}};
    auto it = ids.find(bin);
    if (it == ids.end()) throw std::runtime_error("load_kernel: unknown kernel '" + bin + "'");
    kernel_ = it->second;
  }

  /* blocking, like dpu_launch(DPU_SYNCHRONOUS) */
  void start_kernel() {
    if (kernel_ == FHE_KERNEL_BOOT) return;
    if (meta_.empty()) throw std::runtime_error("start_kernel: no descriptor was copied to \"meta\"");
    check(fhe_launch(ctx_, kernel_, meta_.data(), heap_), "start_kernel");
  }

  void display() {}   /* DPU printf logs have no counterpart */

  void display_memory_status() const {
    std::cout << "Chunk 0:\n";
    chunk_.display_status();
    std::cout << "\n";
  }

  /* `size` in 64-bit words, as PimData passes it; returns {(device index, byte offset)} or {} */
  std::vector<std::pair<size_t, uint32_t>> allocate(size_t size) {
    std::lock_guard<std::mutex> lock(heap_mutex_);
    int64_t off = chunk_.allocate(size * sizeof(uint64_t));
    if (off < 0) {
      std::cerr << "Insufficient memory across all chunks.\n";   /* PimManager.cpp:76-78 */
      return {};
    }
    return {{(size_t)0, (uint32_t)off}};
  }

  void deallocate(const std::vector<std::pair<size_t, uint32_t>> &addrs) {
    std::lock_guard<std::mutex> lock(heap_mutex_);
    for (const auto &a : addrs) chunk_.deallocate(a.second);
  }

  uint32_t getNumDpus() { return 1; }

  /* ---- additions for callers that go past the legacy descriptor path ---- */
  fhe_ctx *context() const { return ctx_; }
  fhe_dptr heap_base() const { return heap_; }
  fhe_dptr device_ptr(uint32_t offset) const { return heap_ + offset; }
  uint32_t requested_dpus() const { return nr_dpus; }
  void sync() { check(fhe_sync(ctx_, -2), "sync"); }

  static void check(int rc, const char *what) {
    if (rc != FHE_OK) throw std::runtime_error(std::string(what) + ": " + fhe_last_error());
  }

private:
  PimManager(uint32_t count, const std::string &profile) : nr_dpus(count) {
    long device = option(profile, "device", 0), heap_mb = option(profile, "heap_mb", 1024);
    if (heap_mb < 1 || heap_mb > 4095) throw std::runtime_error("PimManager: heap_mb must be in [1,4095]");
    uint32_t streams = count < 1 ? 1 : (count > 64 ? 64 : count);
    std::string opts = "streams=" + std::to_string(option(profile, "streams", streams)) +
                       ",arena_mb=" + std::to_string(option(profile, "arena_mb", 256));
    check(fhe_init((int)device, opts.c_str(), &ctx_), "PimManager");
    size_t bytes = (size_t)heap_mb << 20;
    int rc = fhe_alloc(ctx_, bytes, &heap_);
    if (rc != FHE_OK) {
      std::string msg = fhe_last_error();
      fhe_shutdown(ctx_);
      throw std::runtime_error("PimManager: device heap: " + msg);
    }
    chunk_ = DpuMemory(bytes);
    load_kernel(BOOT);
  }

  ~PimManager() {
    fhe_free(ctx_, heap_);
    fhe_shutdown(ctx_);
  }

  PimManager(const PimManager &) = delete;
  PimManager &operator=(const PimManager &) = delete;

  static long option(const std::string &s, const char *key, long dflt) {
    std::string k = std::string(key) + "=";
    size_t p = 0;
    while ((p = s.find(k, p)) != std::string::npos) {
      if (p == 0 || s[p - 1] == ',' || s[p - 1] == ' ') return atol(s.c_str() + p + k.size());
      p += k.size();
    }
    return dflt;
  }

  void bounds(uint32_t offset, size_t bytes, const char *what) const {
    if ((size_t)offset + bytes > chunk_.capacity())
      throw std::runtime_error(std::string(what) + ": range exceeds the device heap");
  }

  inline static PimManager *pim = nullptr;
  inline static std::mutex mutex_;
  std::mutex heap_mutex_;
  fhe_ctx *ctx_ = nullptr;
  fhe_dptr heap_ = 0;
  uint32_t nr_dpus;
  DpuMemory chunk_;
  int kernel_ = FHE_KERNEL_BOOT;
  std::vector<char> meta_;
};

#endif /* _PIM_MANAGER_ */
