// Force-included (-include) ahead of every reference header when building oracle/_ref/plu_ref.
// TEST INFRASTRUCTURE ONLY.  Supplies the std headers the MSVC-era reference forgets on Linux and
// makes the reference's global RNG seedable WITHOUT touching its sources (cmmn.h:240 constructs
// `mt19937(random_device()())`; we rename random_device to a functor that returns $PLU_SEED).
#pragma once
#include <malloc.h>
#include <cstdint>
#include <cfloat>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <cassert>
#include <list>
#include <fstream>
#include <stdexcept>
#include <random>
struct plu_fixed_seed_device {
	typedef unsigned int result_type;
	unsigned int operator()() const {
		const char* s = getenv("PLU_SEED");
		return s ? (unsigned int)strtoul(s, nullptr, 10) : 5489u;
	}
};
#define random_device plu_fixed_seed_device
