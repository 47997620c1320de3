// oracle/shim -- TEST INFRASTRUCTURE: forwards to the mini-cv shim (see ../cvshim.h)
#pragma once
#include "cvshim.h"
