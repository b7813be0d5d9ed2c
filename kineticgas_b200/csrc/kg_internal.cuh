// kg_internal.cuh -- definitions shared by the translation units of libkgcuda.so (not part of the C-ABI).
#pragma once
#include "kg_device.cuh"
#include "../../include/kgcuda.h"
#include <string>
#include <vector>

#define CUDA_OK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
    set_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); return -2; } } while (0)

struct kg_model {
    int device = 0, ncomps = 0, npairs = 0, is_singlecomp = 0, ntpad = 2;
    std::vector<double> mass;
    std::vector<kg_potential_desc> descs;
    std::vector<kg::PairParams> pairs;
    std::vector<double*> d_rows;        // per pair: [KMAX][2+ntpad]
    double** d_rows_by_pair = nullptr;
    kg::PairParams* d_pairs = nullptr;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    std::string err;
    int sm_count = 148;
    // stage timings of the last kg_tcache_create / kg_transport_batch on this model (kg_last_timings)
    double timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int tl_launches = 0;       // kernel launches of the last kg_transfer_lengths_batch
    void set_error(const std::string& s) { err = s; }
};

