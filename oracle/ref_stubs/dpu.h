/* Minimal host-only stand-in for the UPMEM SDK <dpu.h>, just enough for the reference's
 * pim/PimManager.h:13-17,62-110 and src/core/pim/host/PimManager.cpp:20-52 to compile.
 * Our own code (not copied from the SDK): every call is a no-op that reports success;
 * DPU_FOREACH iterates nr_dpus times so PimManager::allocate hands out blocks. */
#ifndef ORACLE_REF_DPU_STUB_H
#define ORACLE_REF_DPU_STUB_H
#include <stdint.h>
#include <stdio.h>
#include <stddef.h>
typedef int dpu_error_t;
#define DPU_OK 0
struct dpu_set_t { uint32_t n; uint32_t idx; };
typedef enum { DPU_XFER_TO_DPU, DPU_XFER_FROM_DPU } dpu_xfer_t;
typedef enum { DPU_XFER_DEFAULT = 0 } dpu_xfer_flags_t;
typedef enum { DPU_SYNCHRONOUS, DPU_ASYNCHRONOUS } dpu_launch_policy_t;
#define DPU_MRAM_HEAP_POINTER_NAME "__sys_used_mram_end"
#define DPU_ASSERT(stmt) do { (void)(stmt); } while (0)
static inline dpu_error_t dpu_alloc(uint32_t n, const char* p, struct dpu_set_t* s) { (void)p; s->n = n; s->idx = 0; return DPU_OK; }
static inline dpu_error_t dpu_free(struct dpu_set_t s) { (void)s; return DPU_OK; }
static inline dpu_error_t dpu_load(struct dpu_set_t s, const char* b, void* x) { (void)s; (void)b; (void)x; return DPU_OK; }
static inline dpu_error_t dpu_launch(struct dpu_set_t s, dpu_launch_policy_t p) { (void)s; (void)p; return DPU_OK; }
static inline dpu_error_t dpu_log_read(struct dpu_set_t s, FILE* f) { (void)s; (void)f; return DPU_OK; }
static inline dpu_error_t dpu_prepare_xfer(struct dpu_set_t s, void* b) { (void)s; (void)b; return DPU_OK; }
static inline dpu_error_t dpu_push_xfer(struct dpu_set_t s, dpu_xfer_t d, const char* sym, uint32_t off, size_t len, dpu_xfer_flags_t f) { (void)s; (void)d; (void)sym; (void)off; (void)len; (void)f; return DPU_OK; }
static inline dpu_error_t dpu_broadcast_to(struct dpu_set_t s, const char* sym, uint32_t off, const void* b, size_t len, dpu_xfer_flags_t f) { (void)s; (void)sym; (void)off; (void)b; (void)len; (void)f; return DPU_OK; }
#define DPU_FOREACH_2(set, dpu) for ((dpu) = (set), (dpu).idx = 0; (dpu).idx < (set).n; (dpu).idx++)
#define DPU_FOREACH_3(set, dpu, i) for ((dpu) = (set), (dpu).idx = 0, (i) = 0; (dpu).idx < (set).n; (dpu).idx++, (i)++)
#define DPU_FOREACH_PICK(_1, _2, _3, NAME, ...) NAME
#define DPU_FOREACH(...) DPU_FOREACH_PICK(__VA_ARGS__, DPU_FOREACH_3, DPU_FOREACH_2)(__VA_ARGS__)
#endif
