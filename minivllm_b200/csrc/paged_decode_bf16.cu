// paged_decode_bf16.cu — decode kernels instantiated for __nv_bfloat16.
#define MVL_T_IS_16BIT 1
#define MVL_T __nv_bfloat16
#include "paged_decode_inst.cuh"
