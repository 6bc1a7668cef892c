// bb_tri_w1.cu -- triangle sweep kernels with W = 1 warps per chain group (see bb_tri_launch.inc).
#define BB_TRI_W 1
#define BB_TRI_FN launch_tri_w1
#include "bb_tri_launch.inc"
