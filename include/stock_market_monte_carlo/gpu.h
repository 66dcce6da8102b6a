// Drop-in for reference include/stock_market_monte_carlo/gpu.h:1-2 (definitions: reference src/gpu.cpp:7-15,
// src/gpu.cu:17-47).  Exported by libsmmc.so.
#pragma once
void vector_add(float *out, float *a, float *b, int n);      // host loop
void vector_add_gpu(float *out, float *a, float *b, int n);  // device kernel behind smmc_vector_add (include/smmc.h)
