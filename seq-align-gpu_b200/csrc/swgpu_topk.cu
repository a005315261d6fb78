// swgpu_topk.cu -- placeholder until the device-side selection lands (next milestone).
#include "../../include/swgpu.h"
extern "C" int swgpu_topk(swgpu_db_t*, int, swgpu_hit_t*, int* n_hits) { if (n_hits) *n_hits = 0; return SWGPU_EINVAL; }
