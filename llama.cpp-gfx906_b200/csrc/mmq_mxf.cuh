// mmq_mxf.cuh — host interface of the native block-scaled MXFP4 kernel (mmq_mxf.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace b200q {

struct MxfArgs {
    int64_t k, m, n;           // dense: dst[n][m] = act[n][k] . w[m][k]^T
    const uint8_t * w;         // block_mxfp4 rows at pitch w_nb1 (experts at w_nb2)
    const float *   act;
    float *         dst;
    long long w_nb1, w_nb2;
    long long a_nb1, a_nb2;    // dense: column stride | MoE: slot, token strides of b
    long long d_nb1, d_nb2;    // dense: column stride | MoE: slot, token strides of dst
    // MoE (ids != nullptr)
    const int32_t * ids;
    long long       ids_nb1;
    int64_t         n_expert, n_used, n_tokens;
    bool            b_bcast;
};

bool   mxf_supported(int type, int64_t k);
size_t mxf_workspace_bytes(int64_t k, int64_t n);
size_t mxf_moe_workspace_bytes(int64_t k, int64_t n_tokens, int64_t n_used, int64_t n_expert);
int    mxf_launch(const MxfArgs & a, void * workspace, cudaStream_t st, int sms, std::string & err);

}  // namespace b200q
