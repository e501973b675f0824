// libnorm_nccl.so -- NCCL all-reduce of the SSIM moments (include/norm_nccl.h).
#include "../../include/norm_nccl.h"

#include <cuda_runtime.h>
#include <nccl.h>

#include <cstring>
#include <new>
#include <string>

struct norm_nccl_comm {
    ncclComm_t comm = nullptr;
    double *dev = nullptr;
    size_t cap = 0;
    cudaStream_t stream = nullptr;
};

namespace {
thread_local std::string g_err;
int32_t fail(const std::string &m) {
    g_err = m;
    return 1;
}
}  // namespace

#define NC(expr)                                                                             \
    do {                                                                                     \
        ncclResult_t r__ = (expr);                                                           \
        if (r__ != ncclSuccess) return fail(std::string(#expr) + ": " + ncclGetErrorString(r__)); \
    } while (0)
#define CU(expr)                                                                             \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) return fail(std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    } while (0)

extern "C" {

int32_t norm_nccl_unique_id(uint8_t id[128]) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    NC(ncclGetUniqueId(&u));
    std::memcpy(id, &u, 128);
    return 0;
}

int32_t norm_nccl_comm_create(const uint8_t id[128], int32_t nranks, int32_t rank, norm_nccl_comm **out) {
    if (!id || !out || nranks < 1 || rank < 0 || rank >= nranks) return fail("bad arguments");
    *out = nullptr;
    norm_nccl_comm *c = new (std::nothrow) norm_nccl_comm();
    if (!c) return fail("out of memory");
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    ncclResult_t r = ncclCommInitRank(&c->comm, nranks, u, rank);
    if (r != ncclSuccess) {
        delete c;
        return fail(std::string("ncclCommInitRank: ") + ncclGetErrorString(r));
    }
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        ncclCommDestroy(c->comm);
        delete c;
        return fail(std::string("cudaStreamCreate: ") + cudaGetErrorString(e));
    }
    *out = c;
    return 0;
}

void norm_nccl_comm_destroy(norm_nccl_comm *c) {
    if (!c) return;
    if (c->dev) cudaFree(c->dev);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->comm) ncclCommDestroy(c->comm);
    delete c;
}

int32_t norm_nccl_allreduce_sum_f64(norm_nccl_comm *c, double *device_buf, size_t count, void *cuda_stream) {
    if (!c || !device_buf || count == 0) return fail("bad arguments");
    NC(ncclAllReduce(device_buf, device_buf, count, ncclDouble, ncclSum, c->comm, static_cast<cudaStream_t>(cuda_stream)));
    return 0;
}

int32_t norm_nccl_allreduce_moments(norm_nccl_comm *c, double *host, size_t count) {
    if (!c || !host || count == 0) return fail("bad arguments");
    if (c->cap < count) {
        if (c->dev) cudaFree(c->dev);
        c->dev = nullptr, c->cap = 0;
        CU(cudaMalloc(&c->dev, count * sizeof(double)));
        c->cap = count;
    }
    CU(cudaMemcpyAsync(c->dev, host, count * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    NC(ncclAllReduce(c->dev, c->dev, count, ncclDouble, ncclSum, c->comm, c->stream));
    CU(cudaMemcpyAsync(host, c->dev, count * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}

int32_t norm_nccl_version(void) {
    int v = 0;
    ncclGetVersion(&v);
    return v;
}

const char *norm_nccl_last_error(void) { return g_err.c_str(); }

}  // extern "C"
