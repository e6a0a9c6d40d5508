// oth_size.cu -- one translation unit per board size: nvcc -DOTH_S=<S>.
// Instantiates every kernel of oth_kernels.cuh for that size and hands the C ABI layer
// (oth_cabi.cu) a table of launchers.
#include "oth_kernels.cuh"

#ifndef OTH_S
#error "compile with -DOTH_S=<board size>"
#endif

#define OTH_CAT2(a, b) a##b
#define OTH_CAT(a, b) OTH_CAT2(a, b)

extern "C" const OthSizeOps *OTH_CAT(oth_size_ops_, OTH_S)() { return SizeLaunch<OTH_S>::ops(); }
