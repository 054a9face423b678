// libqtts_b200.cu — single translation unit of the shared library (kernels are launched from
// the host code below them, so everything is compiled together without relocatable device code).
#include "qtts_kernels.cu"
#include "qtts_batched.cu"
#include "qtts_weights.cu"
#include "qtts_batched_host.cu"
#include "qtts_program.cu"
#include "qtts_api.cu"
#include "qtts_microbench.cu"
