/* C ABI of libtfde_runtime.so -- host-side input pipeline (csrc/runtime/pipeline.cc), no CUDA dependency.
 * Replaces the tf.data C++ runtime behind `from_tensor_slices(...).shuffle(B).repeat().batch(N).prefetch(P)`
 * (ref: mnist_keras_distributed.py:142-145, distributed_with_keras.py:30,54). */
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* features/labels: contiguous arrays of n rows kept alive by the caller. repeat < 0: forever.
 * shuffle_buffer <= 1: no shuffling. depth: prefetch ring size (>= 2). Returns NULL on bad arguments. */
void* tfde_pipe_create(const void* features, int64_t feat_bytes, const void* labels, int64_t label_bytes,
                       int64_t n, int64_t shuffle_buffer, uint64_t seed, int64_t repeat, int64_t batch,
                       int drop_remainder, int depth, int shard_num, int shard_index);
/* Blocks until a batch is ready, copies it into caller memory, returns its row count (0 == end of data). */
int64_t tfde_pipe_next_into(void* pipe, void* feat_dst, void* label_dst);
void tfde_pipe_destroy(void* pipe);
int tfde_runtime_abi_version(void);
#ifdef __cplusplus
}
#endif
