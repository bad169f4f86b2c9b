// Stand-in for the reference's umbrella header src/include/FAST/FAST_directives.hpp (which pulls in
// the FAST framework).  Put this directory first on the include path and the reference's three
// drivers compile UNCHANGED against the B200 library: every `fast::X` they use is nm03::X.
#pragma once
#include <sys/types.h>

#include <iostream>

#include "../../nm03.hpp"

namespace fast = nm03;
