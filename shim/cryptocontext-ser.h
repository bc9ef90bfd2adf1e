// shim/cryptocontext-ser.h -- OpenFHE serialisation header named by src/main.cpp:30-33; everything is in openfhe.h.
#pragma once
#include "openfhe.h"
