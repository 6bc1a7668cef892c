// bb_tri_w2.cu -- triangle sweep kernels with W = 2 warps per chain group (see bb_tri_launch.inc).
#define BB_TRI_W 2
#define BB_TRI_FN launch_tri_w2
#include "bb_tri_launch.inc"
