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
  /* node hand-over (work stealing, include/mvolps_bnb.h): blocking exchange of
   * HOST buffers between the two ranks of a pair.  Send sbytes to rank dst
   * (dst < 0: nothing to send) and receive rbytes from rank src (src < 0:
   * nothing to receive); only the ranks of a pair call it.  May be NULL: the
   * node-sharded tree then runs without hand-over.  0 = ok */
  int (*sendrecv_host)(void *ctx, const void *send, size_t sbytes, int dst, void *recv, size_t rbytes, int src);
  /* all-gather of `bytes` bytes per rank, HOST buffers (recv holds world*bytes, by rank): the
   * result exchange of the round-synchronous tree (MVB_DIST_SYNC, include/mvolps_bnb.h).  May be
   * NULL: that mode is then refused.  0 = ok */
  int (*allgather_host)(void *ctx, const void *send, void *recv, size_t bytes);
} mvb_comm;

#ifdef __cplusplus
}
#endif
#endif
