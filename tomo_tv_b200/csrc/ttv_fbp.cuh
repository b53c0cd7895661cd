// CGLS and FBP/WBP drivers (second tier).  Filled in after the FISTA hot path.
#pragma once
#include "ttv_engine.cuh"
namespace ttv {
inline void cgls_run(Engine&, int) { throw CudaError("CGLS is not implemented yet in this build"); }
inline void fbp_run(Engine&, bool) { throw CudaError("FBP/WBP is not implemented yet in this build"); }
}
