// node_patch.cuh — parameters of patch_nodes (kernels_narrow.cu): fgs_update_nodes for a few rows without a host round trip
#pragma once
#include <cuda_runtime.h>

#include "kernels.cuh"
#include "kernels_narrow.cuh"

namespace fgs {

// a handful of rows travel as kernel arguments, more as a
// compact [row][column] buffer
constexpr int kPatchInline = 4;
constexpr int kPatchMaxCols = NR_FIXED_COLS + 2 * FGS_MAX_SCALARS + FGS_MAX_DEVICES;
struct PatchParams {
    DeriveParams dp;
    const long long* rows; // device [n_dirty][ncols], or nullptr: inline_rows
    const long long* idx;  // device [n_dirty]
    int n_dirty, ncols;
    NodeAnalysis ref;      // the analysis the host decides by
    long long* special;
    int* nn;
    NarrowShifts sh;
    int fit_most, narrow_valid;
    int* d_flag;           // raised when a row leaves `ref`
    volatile int* h_flag;  // the same, in mapped host memory
    long long inline_idx[kPatchInline];
    long long inline_rows[kPatchInline][kPatchMaxCols];
};
cudaError_t launch_patch_nodes(const PatchParams& p, void* stream);


} // namespace fgs
