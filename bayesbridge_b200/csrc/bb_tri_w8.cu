// bb_tri_w8.cu -- triangle sweep kernels with W = 8 warps per chain group (see bb_tri_launch.inc).
#define BB_TRI_W 8
#define BB_TRI_FN launch_tri_w8
#include "bb_tri_launch.inc"
