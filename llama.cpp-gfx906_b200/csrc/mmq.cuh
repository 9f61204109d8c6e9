// mmq.cuh — host interface of the tensor-core (tcgen05) prompt-processing kernel (mmq.cu)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace b200q {

struct MmqArgs {
    int             type;
    int64_t         k, m, n;
    int64_t         ne02, ne03, ne12, ne13;
    long long       w_nb1, w_nb2, w_nb3;
    long long       a_nb1, a_nb2, a_nb3;
    long long       d_nb1, d_nb2, d_nb3;
    const uint8_t * w;
    const float *   act;
    float *         dst;
    bool            reuse_pack = false;  // caller: `act` holds the same values as in the previous call (B200Q_FLAG_ACT_UNCHANGED)
};

void mmq_forget_pack();  // another kernel family is about to overwrite the workspace: no packed image to reuse

struct MmqMoeArgs {  // MUL_MAT_ID through the tensor-core kernel (many tokens)
    int             type;
    int64_t         k, m, n_expert, n_used, n_tokens;
    bool            b_bcast;           // b->ne[1] == 1
    long long       w_nb1, w_nb2;      // expert weights: row, expert
    long long       a_nb1, a_nb2;      // b: slot, token
    long long       ids_nb1;
    long long       d_nb1, d_nb2;      // dst: slot, token
    const uint8_t * w;
    const float *   act;
    const int32_t * ids;
    float *         dst;
};

bool   mmq_supported(int type, int64_t k);
size_t mmq_moe_workspace_bytes(int64_t k, int64_t n_tokens, int64_t n_used, int64_t n_expert);
int    mmq_moe_launch(const MmqMoeArgs & a, void * workspace, cudaStream_t st, int sms, std::string & err);
size_t mmq_workspace_bytes(int64_t k, int64_t n, int64_t nbatch);
void   mmq_set_debug(unsigned long long * dev_buf);  // development aid (clock64 stamps of CTA 0)
int    mmq_launch(const MmqArgs & a, void * workspace, cudaStream_t st, int sms, std::string & err);

}  // namespace b200q
