// bb_tri_w4.cu -- triangle sweep kernels with W = 4 warps per chain group (see bb_tri_launch.inc).
#define BB_TRI_W 4
#define BB_TRI_FN launch_tri_w4
#include "bb_tri_launch.inc"
