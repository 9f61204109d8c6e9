/*
 * ggml-b200.h — entry points exported by the ggml backend plugin libggml-b200.so.
 *
 * These are exactly the symbols the reference's plugin loader binds
 * (ggml/src/ggml-backend-impl.h:209-251 typedefs ggml_backend_init_t / ggml_backend_score_t; looked up with
 * dlsym in ggml/src/ggml-backend-reg.cpp:232-276) plus the named registration function every in-tree backend
 * has (cf. ggml_backend_cuda_reg, ggml/include/ggml-cuda.h).  The vtables behind the returned registry are those
 * of ggml-backend-impl.h:17-207 (API version 1) — see INTEGRATION.md for how a host loads it.
 *
 * Device names are "B2000", "B2001", ... (one per compute-capability 10.x GPU); the registry name is "B200".
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#ifdef __cplusplus
extern "C" {
#endif

struct ggml_backend_reg;

#define GGML_B200_API __attribute__((visibility("default")))

/* registry of the backend; process-lifetime static, thread-safe lazy initialisation */
GGML_B200_API struct ggml_backend_reg * ggml_backend_b200_reg(void);
/* dlopen entry point: returns ggml_backend_b200_reg() */
GGML_B200_API struct ggml_backend_reg * ggml_backend_init(void);
/* 0 when no compute-capability 10.x device is present (the loader then skips the plugin), else 100 */
GGML_B200_API int ggml_backend_score(void);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
