// comm.h — NCCL behind the C-ABI (comm.cu): one process per GPU, SUM reduce of the accumulation buffers at a root.
#ifndef PTB_COMM_H_
#define PTB_COMM_H_
#include <cuda_runtime.h>
#include <stddef.h>
#define COMM_ID_BYTES 128
int comm_unique_id(void *out_id, char *err, size_t cap);
int comm_init(int rank, int nranks, const void *id_bytes, char *err, size_t cap);
int comm_ranks(void);
int comm_rank(void);
int comm_reduce_sum(float *buf, size_t count, int root, cudaStream_t st, char *err, size_t cap);
int comm_broadcast(void *buf, size_t bytes, int root, cudaStream_t st, char *err, size_t cap);
int comm_barrier(cudaStream_t st, char *err, size_t cap);    // all ranks' streams have reached this point; synchronises the stream
int comm_destroy(void);
#endif
