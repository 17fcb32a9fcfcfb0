// pg_tu_reset_e0g1.cu -- reset kernels, variant PG_ENVBOX 0 (obstacle shapes from the kernel parameters), with a general start-pose randomizer (see pg_reset_body.cuh)
#define PG_ENVBOX 0
#define PG_RESET_KERNEL k_reset
#define PG_RESET_HARD_KERNEL k_reset_deferred
#define PG_RESET_GPOSE true
#define PG_RESET_LAUNCH launch_reset_e0g1
#include "pg_reset_variant.inc"
