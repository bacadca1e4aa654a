// Hand-written stand-in for Paracabs' generated configure.hpp
// (reference: dependencies/Paracabs/CMakeLists.txt:84-85, configure.hpp.in).
// TEST INFRASTRUCTURE ONLY (oracle/build_ref.sh).
#pragma once
#define PARACABS_USE_MESSAGE_PASSING  false
#define PARACABS_USE_MPI              false
#define PARACABS_USE_MULTI_THREADING  true
#define PARACABS_USE_OPENMP           true
#define PARACABS_USE_SIMD             false
#define PARACABS_USE_ACCELERATOR      false
#define PARACABS_USE_CUDA             false
#define PARACABS_USE_SYCL             false
