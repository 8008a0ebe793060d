// sample.cuh — launch parameters of the node-sampling stage (sample.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace fgs {

struct SampleParams {
    uint32_t* bitmap;        // [P][words_per_row], truncated in place
    const int64_t* nd;       // derived node table
    const int64_t* pods;     // device pod table
    int64_t* reduce_max;     // [FGS_REDUCE_ROWS][p_pad]
    int64_t* reduce_sum;     // [p_pad] ← len(feasibleNodes) after the walk
    void* walk;              // int2 [P]: {start index, processed nodes} of every pod's walk
    int64_t* next_start;     // device scalar: nextStartNodeIndex after the last pod
    int64_t words_per_row, n_nodes, n_pad, n_pods, p_pad;
    int64_t k;               // numFeasibleNodesToFind
    int64_t start;           // nextStartNodeIndex before the first pod
    void* scratch;           // sample_scratch_bytes(): the block maps of the composed walk (may be NULL: the serial walk runs)
    int64_t scratch_bytes;
};

// the walk over blocks of pods (sample.cu: sample_compose / sample_block_starts): bytes of scratch it needs, 0 = not applicable
int64_t sample_scratch_bytes(int64_t n_nodes, int64_t n_pods);
int sample_launch_count(int64_t n_nodes, int64_t n_pods, bool have_scratch);

int64_t num_feasible_nodes_to_find(int64_t n_nodes, int32_t percentage);
// sample_chain + sample_truncate + reduce_from_bitmap (3 launches)
cudaError_t launch_sampling(const SampleParams& p, void* stream);
// reduce_from_bitmap alone (bitmap, nd, pods, reduce_max, sizes of `p` are read), for the pods with only[p] != 0
cudaError_t launch_reduce_from_bitmap(const SampleParams& p, const int64_t* only, void* stream);

} // namespace fgs
