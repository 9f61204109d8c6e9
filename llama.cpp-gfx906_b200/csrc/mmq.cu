// mmq.cu — placeholder until the tcgen05 kernel lands (n > 8 currently runs as repeated GEMV passes)
#include "mmq.cuh"
namespace b200q {
bool   mmq_supported(int, int64_t) { return false; }
size_t mmq_workspace_bytes(int64_t, int64_t, int64_t) { return 0; }
int    mmq_launch(const MmqArgs &, void *, cudaStream_t, int, std::string & err) { err = "mmq not built"; return -2; }
}  // namespace b200q
