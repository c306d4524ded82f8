/*
 * pim/common.h — host <-> device argument blocks of the intercept surface.
 *
 * Same names, field order and meaning as the reference's wire format
 * (src/core/include/pim/common.h:9-54), so code written against the UPMEM fork compiles
 * unchanged.  `start` is a byte offset inside the device heap (for SCALAR/SCALAR_EQ the second
 * operand's `start` carries the scalar itself, pim/PimData.h:252,271); `size` is in bytes.
 */
#ifndef __PIM_COMMON__
#define __PIM_COMMON__

#include <stdint.h>

#define DIVROUNDUP(x, y) (((x) + (y)-1) / (y))
#define DIV(x, y) ((x) / (y))

struct operands {
  uint64_t start;
  uint32_t size;
};

enum add_sub_mul_kernel {
  SCALAR = 0,
  SCALAR_EQ = 1,
  VECTOR = 2,
  VECTOR_EQ = 3,
};

/* argument block of the modular add / sub / mult kernels */
struct mod_add_sub_mult {
  struct operands op1;
  struct operands op2;
  struct operands res;
  uint64_t mod;
  uint64_t mu;
  enum add_sub_mul_kernel kernel;
};

enum mod_kernel {
  SWITCHMODULUS = 0,
  MOD_NORMAL = 1,
  MODEQ = 2,
  MODBYTWO = 3,
  MODBYTWOEQ = 4,
  MODEXP = 5,
  MODEXPEQ = 6,
  MODINVERSE = 7,
  MODINVERSEEQ = 8,
  MULTIPLYANDROUND = 9,
  MULTIPLYANDROUNDEQ = 10,
  DIVIDEANDROUND = 11,
  DIVIDEANDROUNDEQ = 12
};

struct mod_ops {
  struct operands op;
  struct operands res;
  uint64_t mod;
  uint64_t p;
  uint64_t q;
  enum mod_kernel kernel;
};

#endif /* __PIM_COMMON__ */
