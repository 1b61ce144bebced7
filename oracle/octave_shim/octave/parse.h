// TEST INFRASTRUCTURE ONLY -- see oct.h in this directory.
#include "oct.h"
