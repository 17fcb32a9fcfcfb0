// pg_tu_reset_e1g1.cu -- reset kernels, variant PG_ENVBOX 1 (per-environment cuboid sizes), with a general start-pose randomizer (see pg_reset_body.cuh)
#define PG_ENVBOX 1
#define PG_RESET_KERNEL k_reset_envbox
#define PG_RESET_HARD_KERNEL k_reset_envbox_deferred
#define PG_RESET_GPOSE true
#define PG_RESET_LAUNCH launch_reset_e1g1
#include "pg_reset_variant.inc"
