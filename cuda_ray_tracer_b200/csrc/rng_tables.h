// rng_tables.h — host-built XORWOW skip-ahead tables (rng_tables.cpp), consumed by render.cu.
#pragma once
#include <stdint.h>

#include <vector>

#include "../../include/rt_api.h"

#define RT_RNG_MATRIX_WORDS (160 * 8)  // one 160x160 GF(2) matrix: 160 rows x 8 words (5 used)

void rt_xorwow_seed(uint64_t seed, uint32_t* v5, uint32_t* d);
void rt_xorwow_power_matrices(int count, std::vector<uint32_t>& out);
void rt_xorwow_base_vectors(uint64_t seed, int stride, int count, std::vector<uint32_t>& out);
