// pg_tu_reset_e0g0.cu -- reset kernels, variant PG_ENVBOX 0 (obstacle shapes from the kernel parameters), arm at the stage's reset pose (see pg_reset_body.cuh)
#define PG_ENVBOX 0
#define PG_RESET_KERNEL k_reset
#define PG_RESET_HARD_KERNEL k_reset_deferred
#define PG_RESET_GPOSE false
#define PG_RESET_LAUNCH launch_reset_e0g0
#define PG_OBSERVE_KERNEL k_observe
#define PG_OBSERVE_LAUNCH launch_observe_e0
#define PG_RESET_TRACE_EXPORT
#include "pg_reset_variant.inc"
