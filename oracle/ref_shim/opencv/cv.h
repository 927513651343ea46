// CV shim for the reference build (test infrastructure): everything lives in opencv2/core/core.hpp
#pragma once
#include "../opencv2/core/core.hpp"
