/*
 * pim/DpuMemory.h — host-side bookkeeping of the device heap behind PimManager::allocate.
 *
 * Takes the place of the reference's per-DPU block map (src/core/include/pim/DpuMemory.h:11-82:
 * one 64 MB MRAM chunk per DPU, 4 KB blocks, first fit over a vector<bool>).  Here there is ONE
 * heap per GPU, carved out of HBM by the C ABI (fhe_alloc); this class only hands out byte
 * offsets inside it.  Same public methods (allocate / deallocate / display_status), first fit over
 * an ordered free list with coalescing, 256-byte granularity so every block can be accessed with
 * 128-bit loads.  Not thread safe on its own: PimManager serialises access.
 */
#ifndef _PIM_DPU_MEMORY_
#define _PIM_DPU_MEMORY_

#include "common.h"

#include <cstddef>
#include <cstdint>
#include <iostream>
#include <iterator>
#include <map>

#define PIM_HEAP_ALIGN 256u

class DpuMemory {
public:
  explicit DpuMemory(size_t heap_bytes = 0) : capacity_(heap_bytes) {
    if (heap_bytes) free_[0] = heap_bytes;
  }

  size_t capacity() const { return capacity_; }
  size_t in_use() const { return in_use_; }

  /* returns the byte offset of a block of at least `bytes`, or -1 when the heap is exhausted */
  int64_t allocate(size_t bytes) {
    if (bytes == 0) bytes = PIM_HEAP_ALIGN;
    bytes = DIVROUNDUP(bytes, (size_t)PIM_HEAP_ALIGN) * PIM_HEAP_ALIGN;
    for (auto it = free_.begin(); it != free_.end(); ++it) {
      if (it->second < bytes) continue;
      size_t off = it->first, len = it->second;
      free_.erase(it);
      if (len > bytes) free_[off + bytes] = len - bytes;
      live_[off] = bytes;
      in_use_ += bytes;
      return (int64_t)off;
    }
    return -1;
  }

  void deallocate(uint64_t addr) {
    auto it = live_.find(addr);
    if (it == live_.end()) {
      std::cerr << "Invalid address to deallocate.\n";   /* as DpuMemory.h:39 in the reference */
      return;
    }
    size_t off = it->first, len = it->second;
    live_.erase(it);
    in_use_ -= len;
    auto nx = free_.lower_bound(off);
    if (nx != free_.end() && off + len == nx->first) { len += nx->second; nx = free_.erase(nx); }
    if (nx != free_.begin()) {
      auto pv = std::prev(nx);
      if (pv->first + pv->second == off) { off = pv->first; len += pv->second; free_.erase(pv); }
    }
    free_[off] = len;
  }

  void display_status() const {
    std::cout << "heap " << capacity_ << " B, in use " << in_use_ << " B, " << live_.size()
              << " live block(s), " << free_.size() << " free extent(s)\n";
    for (const auto &kv : live_)
      std::cout << "  [" << kv.first << ", " << kv.first + kv.second << ")\n";
  }

private:
  size_t capacity_ = 0, in_use_ = 0;
  std::map<size_t, size_t> free_;   /* offset -> length */
  std::map<size_t, size_t> live_;
};

#endif /* _PIM_DPU_MEMORY_ */
