// batched_encoder.cuh -- 512-lane encoder under the reference's name (include/core/batched_encoder.cuh:9-49).
#pragma once
#include <cstdint>
#include <cuComplex.h>
#include <cuda_runtime.h>

namespace matrix_fhe {

class BatchedEncoder {
public:
    explicit BatchedEncoder(int n);
    // msg [phi][n*n] complex -> out_re, out_im [w][limb][y][x] (W-eval, X-coeff)
    void encode_to_wntt_eval(const cuDoubleComplex* d_msg_batch, uint64_t* d_out_re, uint64_t* d_out_im, cudaStream_t stream = 0);
    // plain copy of both arrays, as in the reference
    void unpack_eval_p17(const uint64_t* d_in_re, const uint64_t* d_in_im, uint64_t* d_eval_re, uint64_t* d_eval_im,
                         cudaStream_t stream = 0);
    int n() const { return n_; }
    int n2() const { return n2_; }

private:
    int n_;
    int n2_;
    void init_tables_p17();
};

}  // namespace matrix_fhe
