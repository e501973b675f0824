/*
 * norm_nccl.h -- C ABI of libnorm_nccl.so: the one collective of the path.
 *
 * The reference is single-device and has no communication code at all; this is new.  When one
 * giant image is row-band sharded over several GPUs (norm_cuda_chain_band), every stencil
 * stage is local after the halo exchange and the only global step is the SSIM reduction:
 * five f64 moments per frame are summed over ranks with ncclAllReduce before each rank
 * finalises with norm_ssim_from_moments (replacing the whole-image passes of
 * Png::compute_sim, src/png/ssim.rs:18-183).  Frame-sharded batches need no collective.
 *
 * Kept in its own library so that libnorm_cuda.so has no NCCL dependency.
 * One process per GPU; the caller moves the 128-byte unique id from rank 0 to the other
 * ranks by any host channel (torch.distributed store, MPI, a file).
 */
#ifndef NORM_NCCL_H
#define NORM_NCCL_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct norm_nccl_comm norm_nccl_comm;

/* Returns 0 on success, non-zero on error (text from norm_nccl_last_error on this thread). */
int32_t norm_nccl_unique_id(uint8_t id[128]);
/* Collective over all ranks: joins the communicator on the CURRENT CUDA device. */
int32_t norm_nccl_comm_create(const uint8_t id[128], int32_t nranks, int32_t rank, norm_nccl_comm **out);
void norm_nccl_comm_destroy(norm_nccl_comm *comm);
/* In-place sum over ranks of `count` doubles in device memory, on `cuda_stream` (asynchronous). */
int32_t norm_nccl_allreduce_sum_f64(norm_nccl_comm *comm, double *device_buf, size_t count, void *cuda_stream);
/* Convenience for the SSIM moments: host array of `count` doubles in, summed over ranks, out.
 * Uses a small device buffer owned by the communicator; synchronous. */
int32_t norm_nccl_allreduce_moments(norm_nccl_comm *comm, double *host_moments, size_t count);
int32_t norm_nccl_version(void);
const char *norm_nccl_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
