// Resolves the reference's `#include "../configure.hpp"` (src/bindings/pybindings.cpp:1)
// through the -I oracle/ref_cfg/a/b search path. See ../configure.hpp.
#include "../configure.hpp"
