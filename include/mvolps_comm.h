/* mvolps_comm.h -- the communicator vtable shared by the multi-GPU layers
 * (include/mvolps_dist.h: column-sharded tableau; include/mvolps_bnb.h:
 * node-sharded B&B).  Implementations: NCCL inside libglpk_b200.so
 * (mvb_comm_nccl), a single-process loop-back (mvb_comm_local), or callbacks
 * supplied by the caller (torch.distributed/gloo in the CPU tests).
 */
#ifndef MVOLPS_COMM_H
#define MVOLPS_COMM_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mvb_comm {
  void *ctx;
  int rank, world;
  /* all-gather of `bytes` bytes per rank, DEVICE buffers, enqueued on `stream`
   * (a cudaStream_t); recv holds world*bytes.  0 = ok */
  int (*allgather_dev)(void *ctx, const void *send, void *recv, size_t bytes, void *stream);
  /* max-reduction of count doubles in place, HOST buffer (incumbent / bound broadcast) */
  int (*allreduce_max)(void *ctx, double *vals, int count);
  /* release ctx */
  void (*destroy)(void *ctx);
} mvb_comm;

#ifdef __cplusplus
}
#endif
#endif
