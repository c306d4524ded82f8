/*
 * pim/kernel.h — kernel names of the intercept surface (reference: src/core/include/pim/kernel.h:4-8).
 * In the UPMEM fork these are paths of DPU binaries handed to dpu_load; here PimManager::load_kernel
 * maps the same strings to kernel ids of the CUDA backend (FHE_KERNEL_* in fhe_b200.h) — nothing is
 * loaded from disk.
 */
#ifndef __PIM_KERNEL__
#define __PIM_KERNEL__

#define BOOT          "./src/pim/dpu/boot"
#define ELEM_MOD_ADD  "./src/pim/dpu/add-mod"
#define ELEM_MOD_MULT "./src/pim/dpu/mult-mod"
#define ELEM_MOD_SUB  "./src/pim/dpu/sub-mod"
#define ELEM_MOD_OPS  "./src/pim/dpu/mod-ops"

#endif
