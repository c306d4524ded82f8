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
 * Devices.  The reference spreads every vector over `nr_dpus` DPUs: copy_to_pim scatters size / nr_dpus words to each
 * (PimManager.cpp:15-24), allocate reserves DIVROUNDUP(size, nr_dpus) words on each and returns one (chunk index,
 * offset) pair per DPU (PimManager.cpp:56-85), descriptors carry per-DPU sizes (PimData.h:281-290) and start_kernel
 * launches on all of them.  Here the units are GPUs: `gpus=k` in `profile` (or FHE_B200_GPUS=k) takes k devices,
 * `gpus=auto` takes min(nr_dpus, visible GPUs); the default is one (`device=d`).  With k devices getNumDpus() is k,
 * every device has its own heap of `heap_mb`, allocate returns k (device index, byte offset) pairs with equal offsets,
 * copy_to_pim / copy_from_pim scatter / gather size / k words per device (size must be a multiple of k — the reference
 * drops the remainder, here it is an error), and start_kernel runs the descriptor on every device at once
 * (fhe_group_launch).  `nr_dpus` also sizes the stream pool of each device (clamped to [1,64]).  `profile` is the
 * backend option string ("device=0,gpus=1,heap_mb=1024,streams=4,arena_mb=256").  There is no CPU fallback: without a
 * B200 the constructor throws.  Header only; link with libfhe_b200.so.
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
    case 0: {
      const uint32_t per = split(size, "copy_to_pim");
      bounds(offset, (size_t)per * 8, "copy_to_pim");
      for (size_t i = 0; i < heaps_.size(); i++)
        check(fhe_h2d(fhe_group_ctx(group_, (int)i), heaps_[i] + offset, (const uint64_t *)buf + i * per, (size_t)per * 8, 0),
              "copy_to_pim");
      for (size_t i = 0; i < heaps_.size(); i++) check(fhe_sync(fhe_group_ctx(group_, (int)i), 0), "copy_to_pim");
      break;
    }
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
    const uint32_t per = split(size, "copy_from_pim");
    bounds(offset, (size_t)per * 8, "copy_from_pim");
    for (size_t i = 0; i < heaps_.size(); i++)
      check(fhe_d2h(fhe_group_ctx(group_, (int)i), buf + i * per, heaps_[i] + offset, (size_t)per * 8, 0), "copy_from_pim");
    for (size_t i = 0; i < heaps_.size(); i++) check(fhe_sync(fhe_group_ctx(group_, (int)i), 0), "copy_from_pim");
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
    check(fhe_group_launch(group_, kernel_, meta_.data(), heaps_.data()), "start_kernel");
  }

  void display() {}   /* DPU printf logs have no counterpart */

  void display_memory_status() const {
    for (size_t i = 0; i < chunks_.size(); i++) {
      std::cout << "Chunk " << i << ":\n";
      chunks_[i].display_status();
      std::cout << "\n";
    }
  }

  /* `size` in 64-bit words, as PimData passes it: DIVROUNDUP(size, devices) words on every device (PimManager.cpp:58-59);
   * returns one (device index, byte offset) pair per device — the offsets are equal, the descriptor route addresses
   * every device with the first one (PimData.h:281) — or {} when a heap is exhausted */
  std::vector<std::pair<size_t, uint32_t>> allocate(size_t size) {
    std::lock_guard<std::mutex> lock(heap_mutex_);
    const size_t per = DIVROUNDUP(size, chunks_.size());
    std::vector<std::pair<size_t, uint32_t>> out;
    for (size_t i = 0; i < chunks_.size(); i++) {
      int64_t off = chunks_[i].allocate(per * sizeof(uint64_t));
      if (off < 0 || (!out.empty() && (uint32_t)off != out[0].second)) {
        std::cerr << (off < 0 ? "Insufficient memory across all chunks.\n" : "Allocation failed in one of the chunks.\n");
        if (off >= 0) chunks_[i].deallocate((uint64_t)off);
        for (const auto &a : out) chunks_[a.first].deallocate(a.second);        /* roll back (PimManager.cpp:69-72) */
        return {};
      }
      out.push_back({i, (uint32_t)off});
    }
    return out;
  }

  void deallocate(const std::vector<std::pair<size_t, uint32_t>> &addrs) {
    std::lock_guard<std::mutex> lock(heap_mutex_);
    for (const auto &a : addrs)
      if (a.first < chunks_.size()) chunks_[a.first].deallocate(a.second);
  }

  uint32_t getNumDpus() { return (uint32_t)heaps_.size(); }

  /* ---- additions for callers that go past the legacy descriptor path ---- */
  fhe_group *group() const { return group_; }
  uint32_t devices() const { return (uint32_t)heaps_.size(); }
  fhe_ctx *context(uint32_t rank = 0) const { return fhe_group_ctx(group_, (int)rank); }
  fhe_dptr heap_base(uint32_t rank = 0) const { return heaps_[rank]; }
  fhe_dptr device_ptr(uint32_t offset, uint32_t rank = 0) const { return heaps_[rank] + offset; }
  uint32_t requested_dpus() const { return nr_dpus; }
  void sync() { check(fhe_group_sync(group_), "sync"); }

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
    /* how many GPUs: profile `gpus=k|auto`, else FHE_B200_GPUS, else one (the device named by `device=`) */
    std::string want = text_option(profile, "gpus");
    if (want.empty() && getenv("FHE_B200_GPUS")) want = getenv("FHE_B200_GPUS");
    std::vector<int> devs;
    if (want.empty() || want == "1") {
      devs.push_back((int)device);
      check(fhe_group_create(1, devs.data(), opts.c_str(), &group_), "PimManager");
    } else {
      int k = want == "auto" ? -(int)(count < 1 ? 1 : count) : atoi(want.c_str());
      if (k == 0) throw std::runtime_error("PimManager: gpus must be a count or 'auto'");
      if (k < 0) {                                    /* auto: min(nr_dpus, visible) — probe with a group of everything */
        fhe_group *all = nullptr;
        check(fhe_group_create(0, nullptr, opts.c_str(), &all), "PimManager");
        int have = fhe_group_size(all);
        fhe_group_destroy(all);
        k = -k < have ? -k : have;
      }
      check(fhe_group_create(k, nullptr, opts.c_str(), &group_), "PimManager");
    }
    size_t bytes = (size_t)heap_mb << 20;
    for (int i = 0; i < fhe_group_size(group_); i++) {
      fhe_dptr h = 0;
      int rc = fhe_alloc(fhe_group_ctx(group_, i), bytes, &h);
      if (rc != FHE_OK) {
        std::string msg = fhe_last_error();
        fhe_group_destroy(group_);
        throw std::runtime_error("PimManager: device heap: " + msg);
      }
      heaps_.push_back(h);
      chunks_.emplace_back(bytes);
    }
    load_kernel(BOOT);
  }

  ~PimManager() {
    for (size_t i = 0; i < heaps_.size(); i++) fhe_free(fhe_group_ctx(group_, (int)i), heaps_[i]);
    fhe_group_destroy(group_);
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

  static std::string text_option(const std::string &s, const char *key) {
    std::string k = std::string(key) + "=";
    size_t p = 0;
    while ((p = s.find(k, p)) != std::string::npos) {
      if (p == 0 || s[p - 1] == ',' || s[p - 1] == ' ') {
        size_t e = s.find_first_of(", ", p + k.size());
        return s.substr(p + k.size(), e == std::string::npos ? std::string::npos : e - p - k.size());
      }
      p += k.size();
    }
    return "";
  }

  void bounds(uint32_t offset, size_t bytes, const char *what) const {
    if ((size_t)offset + bytes > chunks_[0].capacity())
      throw std::runtime_error(std::string(what) + ": range exceeds the device heap");
  }

  /* words per device of a `size`-word transfer: size / devices, as PimManager.cpp:17 — a remainder is refused */
  uint32_t split(uint32_t size, const char *what) const {
    const uint32_t k = (uint32_t)heaps_.size();
    if (size % k) throw std::runtime_error(std::string(what) + ": " + std::to_string(size) + " words do not divide over " +
                                           std::to_string(k) + " devices");
    return DIV(size, k);
  }

  inline static PimManager *pim = nullptr;
  inline static std::mutex mutex_;
  std::mutex heap_mutex_;
  fhe_group *group_ = nullptr;
  std::vector<fhe_dptr> heaps_;        /* one heap per device */
  uint32_t nr_dpus;
  std::vector<DpuMemory> chunks_;      /* its bookkeeping (DpuMemory per chunk, as PimManager.h:113-118) */
  int kernel_ = FHE_KERNEL_BOOT;
  std::vector<char> meta_;
};

#endif /* _PIM_MANAGER_ */
