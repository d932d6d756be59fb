// paged_decode_f16.cu — decode kernels instantiated for __half.
#define MVL_T_IS_16BIT 1
#define MVL_T __half
#include "paged_decode_inst.cuh"
