// paged_decode_f32.cu — decode kernels instantiated for float.
#define MVL_T float
#include "paged_decode_inst.cuh"
