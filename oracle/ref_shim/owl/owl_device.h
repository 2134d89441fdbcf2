// Stand-in for the un-vendored OWL / OptiX header of the same name.
// Test infrastructure only: lets the reference device TU compile as host C++.
#pragma once
#include "owl_shim.h"
