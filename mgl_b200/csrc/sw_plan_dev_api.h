/*
 * sw_plan_dev_api.h -- what the engine sees of the device planner (sw_plan_dev.cu).
 */
#ifndef MGL_B200_SW_PLAN_DEV_API_H
#define MGL_B200_SW_PLAN_DEV_API_H
#include <cuda_runtime.h>

#include "sw_plan_dev.h"

namespace mglsw {

/* carve-up of the planner's work buffer for a capacity of n pairs (byte offsets; computed once per capacity) */
struct PlanDevLayout {
    size_t key0, key1, idx0, idx1, uhead, urank, u3, u3x, state, cnt, btot, wtot, wtot3, bar, total;
    u32 n_cap, ucap;
};

/* unit descriptors a chunk of n pairs can need: one per pair plus the padding of every launch's last warp */
u32 plan_dev_unit_capacity(u32 n);
PlanDevLayout plan_dev_layout(u32 n_cap);

/*
 * Plans one chunk with ONE kernel launch (16 blocks) on `st`: `in` holds DEVICE pointers to the chunk's pair table; `units`
 * (plan_dev_unit_capacity(n) descriptors) receives what the fill and traceback kernels read; `work` is lay.total
 * bytes; h_summary (pinned) is complete when the stream reaches the end of this call's work.
 */
cudaError_t plan_dev_run(const PlanDevIn &in, const PlanDevLayout &lay, UnitDesc *units, void *work, PlanSummary *h_summary, cudaStream_t st);

}  // namespace mglsw
#endif
