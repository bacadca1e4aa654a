// Hand-written stand-in for the header CMake would generate from the reference's
// src/configure.hpp.in (reference: CMakeLists.txt:83). TEST INFRASTRUCTURE ONLY:
// used by oracle/build_ref.sh to compile the unmodified reference sources where they
// lie under /root/reference into oracle/_ref/. Not product code.
#pragma once
#define MAGRITTE_FOLDER        "/root/repo/oracle/ref_home"
#define MAGRITTE_VERSION       "0.9.3"
#define PYTHON_IO               true
#define MPI_PARALLEL            false
#define OMP_PARALLEL            true
#define GPU_ACCELERATION        false
