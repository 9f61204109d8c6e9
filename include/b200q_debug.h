/* b200q_debug.h — development aids exported by libb200q.so.  NOT part of the drop-in boundary (include/b200q.h): nothing
 * here has a counterpart in the reference, and the plugin never calls any of it.  The host-only plan probes are what
 * tests/test_plan_cpu.py checks the launch planners through (no GPU needed); the stamp buffers feed tools/quick_bench.py,
 * tools/mmq_ts.py and tools/list_timeline.py. */
#ifndef B200Q_DEBUG_H
#define B200Q_DEBUG_H

#include "b200q.h"

#ifdef __cplusplus
extern "C" {
#endif

/* device buffer of [grid][8] u64 that the GEMV CTAs fill with %globaltimer stamps (NULL = off) */
B200Q_API void b200q_debug_timestamps(void * dev_buf);
/* host-mapped buffer that stuck waiters of gemv_kq report into; a no-op unless the library was built with
 * -DB200Q_DEBUG_HANG (the release kernels carry no hang reporting) */
B200Q_API void b200q_debug_hang_buffer(void * host_mapped);
/* device buffer the tensor-core kernel's development build (B200Q_MMQ_DEV) stamps per k-chunk (NULL = off) */
B200Q_API void b200q_debug_mmq_stamps(void * dev_buf);
/* device buffer the list kernel's trace build (-DB200Q_LIST_TRACE) stamps per generation (NULL = off) */
B200Q_API void b200q_debug_list_stamps(void * dev_buf);
/* host only: the launch plan gemv_mma would get on a device with `sms` SMs and `max_smem` bytes of shared memory per CTA.
 * out[16] = nwarps, cluster, kc, chunks, kp, cpm, pitch, stage_bytes, nstages, groups, grid, smem, col_pitch, off_stage,
 * chunk_bytes, row_bytes.  Returns 1 if a plan exists. */
B200Q_API int b200q_debug_plan_gemv_mma(int type, int64_t m, int64_t k, int n, int sms, int max_smem, int * out);
/* host only: the plan of a matmul list (launch cuts, generations, barrier points, ring geometry) */
B200Q_API int b200q_debug_plan_list(const int * types, const int64_t * m, const int64_t * k, const int * dep, const int * same_act,
                                    int n_ops, int world, int sms, int max_smem, int * out, int * seg_out, int max_segs);

#ifdef __cplusplus
}
#endif
#endif
