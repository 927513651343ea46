// CV shim for the reference build (test infrastructure): everything lives in core/core.hpp
#pragma once
#include "../core/core.hpp"
