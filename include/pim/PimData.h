/*
 * pim/PimData.h — device-vector handle of the intercept surface, over the B200 C ABI.
 *
 * Same class name, constructors, operators and named methods as the reference
 * (src/core/include/pim/PimData.h:14-335).  Every operation takes the reference's own route:
 * fill the descriptor of pim/common.h -> PimManager::load_kernel -> copy_to_pim(..., 1, "meta") ->
 * start_kernel (PimData.h:275-305, 307-329 there); behind start_kernel sits fhe_launch and the
 * sm_100a element-wise kernels instead of a DPU binary.  Operands stay resident in HBM between
 * operations; the host copy is refreshed only by sync().
 *
 * Deliberate differences from the reference (SURVEY §8(b) "Ownership"/"Errors"):
 *   - the device block is reference counted, so the shallow operator= (PimData.h:36-44 there) no
 *     longer leads to a double free, and a default-constructed PimData can be destroyed;
 *   - failures throw std::runtime_error instead of aborting (DPU_ASSERT) or indexing an empty
 *     allocation (PimData.h:19-20 there);
 *   - nothing is printed per operation (the reference prints operand offsets, PimData.h:291-292).
 * The operators without a modulus keep the reference's default mod = UINT64_MAX: + and - then wrap
 * exactly like NativeIntegerT::ModAddFastEq / ModSubFastEq do with that modulus; * requires a
 * modulus below 2^61 and throws otherwise (the reference's DPU result for it is meaningless).
 */
#ifndef _PIM_DATA_
#define _PIM_DATA_

#include "PimManager.h"
#include "kernel.h"

#include <cstdint>
#include <initializer_list>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

class PimData {
public:
  PimData(std::initializer_list<uint64_t> data) : host_data(data) {
    pim = PimManager::getPim(4);
    size = (uint32_t)data.size();
    acquire();
    pim->copy_to_pim(host_data.data(), size, offset());
  }

  /* not in the reference: construct from a host vector */
  explicit PimData(const std::vector<uint64_t> &data) : host_data(data) {
    pim = PimManager::getPim(4);
    size = (uint32_t)data.size();
    acquire();
    pim->copy_to_pim(host_data.data(), size, offset());
  }

  PimData(uint64_t size_set) : size((uint32_t)size_set) {
    pim = PimManager::getPim(1);
    acquire();
  }

  PimData() {}

  ~PimData() {}   /* the block is released when its last owner goes away */

  /* device -> host_data */
  void sync() {
    if (!block) return;
    host_data.resize(size);
    pim->copy_from_pim(host_data.data(), size, offset());
  }
  const std::vector<uint64_t> &host() const { return host_data; }
  uint32_t length() const { return size; }
  fhe_dptr device_ptr() const { return block ? pim->device_ptr(offset()) : 0; }

  PimData &operator=(const PimData &rhs) {
    if (this != &rhs) {
      size = rhs.size;
      pim = rhs.pim;
      block = rhs.block;            /* shared, counted */
      metadata = rhs.metadata;
    }
    return *this;
  }
  PimData(const PimData &rhs) { *this = rhs; }

  PimData &operator+=(const uint64_t &rhs) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_ADD); return *this; }
  PimData &operator+=(const PimData &rhs) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_ADD); return *this; }
  PimData operator+(const PimData &rhs) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_ADD); return ret; }
  PimData operator+(const uint64_t &rhs) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_ADD); return ret; }

  PimData &operator-=(const uint64_t &rhs) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_SUB); return *this; }
  PimData &operator-=(const PimData &rhs) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_SUB); return *this; }
  PimData operator-(const PimData &rhs) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_SUB); return ret; }
  PimData operator-(const uint64_t &rhs) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_SUB); return ret; }

  PimData &operator*=(const uint64_t &rhs) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_MULT); return *this; }
  PimData &operator*=(const PimData &rhs) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_MULT); return *this; }
  PimData operator*(const PimData &rhs) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_MULT); return ret; }
  PimData operator*(const uint64_t &rhs) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_MULT); return ret; }

  /* explicit modulus (PimData.h:112-168 in the reference) */
  PimData ModAdd(const uint64_t &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_ADD, modulus); return ret; }
  PimData ModAdd(const PimData &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_ADD, modulus); return ret; }
  void ModAddEq(const uint64_t &rhs, const uint64_t &modulus) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_ADD, modulus); }
  void ModAddEq(const PimData &rhs, const uint64_t &modulus) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_ADD, modulus); }

  PimData ModSub(const uint64_t &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_SUB, modulus); return ret; }
  PimData ModSub(const PimData &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_SUB, modulus); return ret; }
  void ModSubEq(const uint64_t &rhs, const uint64_t &modulus) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_SUB, modulus); }
  void ModSubEq(const PimData &rhs, const uint64_t &modulus) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_SUB, modulus); }

  PimData ModMul(const uint64_t &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, SCALAR, ELEM_MOD_MULT, modulus); return ret; }
  PimData ModMul(const PimData &rhs, const uint64_t &modulus) { PimData ret(size); do_ops(rhs, ret, VECTOR, ELEM_MOD_MULT, modulus); return ret; }
  void ModMulEq(const uint64_t &rhs, const uint64_t &modulus) { do_ops(rhs, SCALAR_EQ, ELEM_MOD_MULT, modulus); }
  void ModMulEq(const PimData &rhs, const uint64_t &modulus) { do_ops(rhs, VECTOR_EQ, ELEM_MOD_MULT, modulus); }
  void MultWithOutMod(const PimData &) {}   /* empty in the reference too (PimData.h:169) */

  /* mod operations (PimData.h:174-233 in the reference; CPU definitions mubintvecnat.cpp:111-195,
   * 370-476).  `om` / `modulus` is the vector's current modulus. */
  PimData Mod(const uint64_t &om, const uint64_t &nm) { PimData ret(size); do_mod_ops(MOD_NORMAL, ret.offset(), ret.size, om, nm, 0); return ret; }
  void ModEq(const uint64_t &om, const uint64_t &nm) { do_mod_ops(MODEQ, 0, 0, om, nm, 0); }
  void switchmod(const uint64_t &om, const uint64_t &nm) { do_mod_ops(SWITCHMODULUS, 0, 0, om, nm, 0); }

  PimData ModByTwo(const uint64_t &modulus) { PimData ret(size); do_mod_ops(MODBYTWO, ret.offset(), ret.size, modulus, 0, 0); return ret; }
  void ModByTwoEq(const uint64_t &modulus) { do_mod_ops(MODBYTWOEQ, 0, 0, modulus, 0, 0); }

  PimData ModExp(const uint64_t &modulus, const uint64_t b) { PimData ret(size); do_mod_ops(MODEXP, ret.offset(), ret.size, modulus, b, 0); return ret; }
  void ModExpEq(const uint64_t &modulus, const uint64_t b) { do_mod_ops(MODEXPEQ, 0, 0, modulus, b, 0); }

  PimData ModInverse(uint64_t modulus) { PimData ret(size); do_mod_ops(MODINVERSE, ret.offset(), ret.size, modulus, 0, 0); return ret; }
  void ModInverseEq(uint64_t modulus) { do_mod_ops(MODINVERSEEQ, 0, 0, modulus, 0, 0); }

  PimData MultiplyAndRound(const uint64_t &modulus, const uint64_t &p, const uint64_t &q) {
    PimData ret(size);
    do_mod_ops(MULTIPLYANDROUND, ret.offset(), ret.size, modulus, p, q);
    return ret;
  }
  void MultiplyAndRoundEq(const uint64_t &modulus, const uint64_t &p, const uint64_t &q) { do_mod_ops(MULTIPLYANDROUNDEQ, 0, 0, modulus, p, q); }

  PimData DivideAndRound(const uint64_t &modulus, const uint64_t &q) { PimData ret(size); do_mod_ops(DIVIDEANDROUND, ret.offset(), ret.size, modulus, q, 0); return ret; }
  void DivideAndRoundEq(const uint64_t &modulus, const uint64_t &q) { do_mod_ops(DIVIDEANDROUNDEQ, 0, 0, modulus, q, 0); }

private:
  struct Block {   /* one device allocation, freed when the last PimData sharing it dies */
    PimManager *pim = nullptr;
    std::vector<std::pair<size_t, uint32_t>> addrs;
    Block() = default;
    Block(const Block &) = delete;
    Block &operator=(const Block &) = delete;
    ~Block() { if (pim && !addrs.empty()) pim->deallocate(addrs); }
  };

  void acquire() {
    metadata = pim->allocate(size);
    if (metadata.empty()) throw std::runtime_error("PimData: device heap exhausted");
    block = std::make_shared<Block>();
    block->pim = pim;
    block->addrs = metadata;
  }

  uint32_t offset() const {
    if (metadata.empty()) throw std::runtime_error("PimData: no device allocation");
    return metadata[0].second;
  }

  static constexpr uint64_t NO_MOD = std::numeric_limits<uint64_t>::max();

  void do_ops(const PimData &rhs, const PimData &res, enum add_sub_mul_kernel kernel, const std::string &ops,
              uint64_t mod = NO_MOD, uint64_t mu = 0) {
    if (rhs.size != size) throw std::runtime_error("PimData: operands of different length");
    common_do_ops(offset(), size * sizeof(uint64_t), rhs.offset(), rhs.size * sizeof(uint64_t), res.offset(),
                  res.size * sizeof(uint64_t), kernel, ops, mod, mu);
  }
  void do_ops(const uint64_t &rhs, const PimData &res, enum add_sub_mul_kernel kernel, const std::string &ops,
              uint64_t mod = NO_MOD, uint64_t mu = 0) {
    common_do_ops(offset(), size * sizeof(uint64_t), rhs, 0, res.offset(), res.size * sizeof(uint64_t), kernel, ops,
                  mod, mu);
  }
  void do_ops(const PimData &rhs, enum add_sub_mul_kernel kernel, const std::string &ops, uint64_t mod = NO_MOD,
              uint64_t mu = 0) {
    if (rhs.size != size) throw std::runtime_error("PimData: operands of different length");
    common_do_ops(offset(), size * sizeof(uint64_t), rhs.offset(), rhs.size * sizeof(uint64_t), 0, 0, kernel, ops,
                  mod, mu);
  }
  void do_ops(const uint64_t &rhs, enum add_sub_mul_kernel kernel, const std::string &ops, uint64_t mod = NO_MOD,
              uint64_t mu = 0) {
    common_do_ops(offset(), size * sizeof(uint64_t), rhs, 0, 0, 0, kernel, ops, mod, mu);
  }

  void common_do_ops(uint64_t op1_start, uint64_t op1_size, uint64_t op2_start, uint64_t op2_size,
                     uint64_t res_start, uint64_t res_size, enum add_sub_mul_kernel kernel, const std::string &ops,
                     uint64_t mod, uint64_t mu) {
    struct mod_add_sub_mult meta;
    meta.op1.start = op1_start;
    meta.op1.size  = (uint32_t)DIV(op1_size, pim->getNumDpus());
    meta.op2.start = op2_start;
    meta.op2.size  = (uint32_t)DIV(op2_size, pim->getNumDpus());
    meta.mod       = mod;
    meta.mu        = mu;           /* recomputed on the device side; kept for wire compatibility */
    meta.res.start = res_start;
    meta.res.size  = (uint32_t)DIV(res_size, pim->getNumDpus());
    meta.kernel    = kernel;
    pim->load_kernel(ops);
    pim->copy_to_pim(&meta, sizeof(meta), 0, 1, "meta");
    pim->start_kernel();
    pim->display();
  }

  void do_mod_ops(enum mod_kernel kernel, uint32_t res_start, uint32_t res_size, uint64_t mod, uint64_t p,
                  uint64_t q) {
    struct mod_ops ops;
    ops.kernel    = kernel;
    ops.op.start  = offset();
    ops.op.size   = (uint32_t)DIV(size * sizeof(uint64_t), pim->getNumDpus());
    ops.res.start = res_start;
    ops.res.size  = (uint32_t)DIV(res_size * sizeof(uint64_t), pim->getNumDpus());
    ops.mod       = mod;
    ops.p         = p;
    ops.q         = q;
    pim->load_kernel(ELEM_MOD_OPS);
    pim->copy_to_pim(&ops, sizeof(ops), 0, 1, "meta");
    pim->start_kernel();
    pim->display();
  }

  std::vector<uint64_t> host_data;
  uint32_t size = 0;
  PimManager *pim = nullptr;
  std::shared_ptr<Block> block;
  std::vector<std::pair<size_t, uint32_t>> metadata;
};

#endif /* _PIM_DATA_ */
