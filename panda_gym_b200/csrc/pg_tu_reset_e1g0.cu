// pg_tu_reset_e1g0.cu -- reset kernels, variant PG_ENVBOX 1 (per-environment cuboid sizes), arm at the stage's reset pose (see pg_reset_body.cuh)
#define PG_ENVBOX 1
#define PG_RESET_KERNEL k_reset_envbox
#define PG_RESET_HARD_KERNEL k_reset_envbox_deferred
#define PG_RESET_GPOSE false
#define PG_RESET_LAUNCH launch_reset_e1g0
#define PG_OBSERVE_KERNEL k_observe_envbox
#define PG_OBSERVE_LAUNCH launch_observe_e1
#include "pg_reset_variant.inc"
