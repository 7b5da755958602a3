// Glue shared by the extern "C" entry points: error capture and launch checks.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>

#include "../../include/lbmff.h"
#include "common.cuh"

namespace lbmff {

void set_last_error(const char* where, cudaError_t e);
void set_last_error_msg(const char* where, const char* msg);

inline int check_launch(const char* where) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error(where, e);
    return (int)e;
  }
  return 0;
}

inline int bad_arg(const char* where, const char* msg) {
  set_last_error_msg(where, msg);
  return (int)cudaErrorInvalidValue;
}

}  // namespace lbmff

#define LBMFF_REQUIRE(cond, msg) \
  if (!(cond)) return lbmff::bad_arg(__func__, msg)
