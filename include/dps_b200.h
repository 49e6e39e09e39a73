/* dps_b200 native library -- C ABI.
 *
 * All entry points are plain C functions loaded with ctypes (dps_b200/ops/_lib.py); pointers are raw
 * device addresses (torch `data_ptr()`), `stream` is a cudaStream_t.  Every function returns 0 on
 * success or a non-zero code, with the message available from dps_last_error().
 *
 * "reference:" notes name the part of Jjjing2023/Distributed-Parameter-Server-for-ML-Training that the
 * entry point replaces (paths relative to the reference root).
 */
#ifndef DPS_B200_H
#define DPS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPS_MAX_RANKS 32   /* reference validates 1..32 workers (src/parameter_server/server.py:424) */
#define DPS_MAX_SLOTS 64
#define DPS_MAX_PLAN 16
#define DPS_HIST_BINS 16

/* Layout of the per-rank flag block (uint32 words, peer-mapped, 4 KiB). */
#define DPS_F_ARRIVE 0          /* [32] sync push: arrivals per source rank                     */
#define DPS_F_PULLCOUNT 32      /* pull deliveries received (all owners)                         */
#define DPS_F_ERROR 33          /* sticky error bits (DPS_ERR_*)                                 */
#define DPS_F_MISSING 34        /* mask of ranks missing at the last timed-out barrier           */
#define DPS_F_TORN_PULLS 35     /* async pulls that stayed inconsistent after retries            */
#define DPS_F_DROPPED 36        /* async pushes skipped because the owner's ring had no room     */
#define DPS_F_ALIVE 37          /* sync: mask of source ranks this owner still waits for (set to all W at start) */
#define DPS_F_UPDATE_NS 40       /* [2] 64-bit: nanoseconds spent in the reduce/apply loops (CTA 0), device-timed */
#define DPS_F_ROUND 39          /* sync round this owner's reduce last entered (read by the gRPC door) */
#define DPS_F_PULLEXPECT 38     /* pull deliveries expected so far (accumulated per round: live owners x CTAs)   */
#define DPS_F_HEARTBEAT 64      /* [32] last step seen from each source rank                     */
#define DPS_F_RING_TAIL 96      /* async ring: next ticket (peer atomics)                        */
#define DPS_F_RING_HEAD 97      /* async ring: tickets consumed                                  */
#define DPS_F_VERSION 98        /* shard version counter (== global_step of this shard)          */
#define DPS_F_PUBSEQ 99         /* seq-lock of the published bf16 shard (odd = being rewritten)  */
#define DPS_F_APPLY_DONE 100    /* local CTA counter of async_apply                              */
#define DPS_F_SLOTS 128         /* [64][4] {done, src, version_seen, pad}                        */
#define DPS_F_STAT_ACCEPTED 512
#define DPS_F_STAT_REJECTED 513
#define DPS_F_STAT_SUM 514
#define DPS_F_STAT_MAX 515
#define DPS_F_STAT_HIST 520     /* [16]                                                          */
#define DPS_F_STAT_LAST 544     /* [32] last staleness per source (reference keeps only this)    */
#define DPS_F_PLAN 640          /* n, new_version, new_head, slots[16], weights[16]              */
#define DPS_F_TICKET 704        /* [32][2] local {tag, ticket} per owner                         */
#define DPS_F_SEEN 800          /* [32] version of each shard at this worker's last pull         */
#define DPS_F_DROP 832          /* [32] local: this step's push to owner o was skipped            */
#define DPS_F_SEGCOUNT 864      /* [16] segmented pull: deliveries received per layer group           */
#define DPS_F_SEGEXPECT 880     /* [16] segmented pull: deliveries expected so far per layer group    */
#define DPS_F_CONSUMED 896      /* [32] pushes of extra source x (rank W+x) folded into a round so far     */
#define DPS_F_CONSUMED_ROUND 928 /* [32] round that consumed the latest push of extra source x           */
#define DPS_F_INCLUDE 960       /* mask of extra sources folded into the CURRENT round (remote_gate_kernel) */
#define DPS_MAX_SEGS 16
#define DPS_FLAG_WORDS 1024

#define DPS_ERR_BARRIER_TIMEOUT 1u
#define DPS_ERR_PULL_TIMEOUT 2u
#define DPS_ERR_RING_FULL 4u

/* Pointer tables of one job, indexed by rank.  Entries of other ranks are peer mappings (CUDA IPC);
 * buffers that are only touched by their owner (grad, master, mom) may be NULL for remote ranks.
 * rank_lo/rank_hi: ranks driven by this launch (production: [rank, rank+1); the single-GPU emulation
 * used by the tests drives all ranks from one launch, see tests/test_ps_kernels.py). */
typedef struct DpsTables {
  int32_t world;
  int32_t rank_lo;
  int32_t rank_hi;
  int32_t block_elems;    /* BS: block-cyclic sharding granule                         */
  int32_t ring_slots;     /* R: async ring depth per owner                             */
  int32_t n_extra;        /* extra source slots after the W on-box workers (remote workers of the gRPC door) */
  int64_t shard_elems;    /* E: elements per shard (multiple of BS)                    */
  int64_t n_small;        /* arena prefix also delivered in fp32 (BN affine, fc bias)  */
  float* grad[DPS_MAX_RANKS];            /* worker: fp32 gradient arena [W*E]                    */
  float* master[DPS_MAX_RANKS];          /* owner: fp32 master shard [E]                         */
  float* mom[DPS_MAX_RANKS];             /* owner: momentum shard [E]                            */
  uint16_t* stage[DPS_MAX_RANKS];            /* owner: sync staging [W][E] fp16 (or fp32)            */
  uint16_t* ring[DPS_MAX_RANKS];             /* owner: async ring [R][E] fp16                        */
  uint16_t* wbf16[DPS_MAX_RANKS];            /* worker: bf16 weight arena [W*E]                      */
  float* small[DPS_MAX_RANKS];           /* worker: fp32 copy of arena[0:n_small]                */
  uint16_t* pub[DPS_MAX_RANKS];              /* owner: published bf16 shard [E] (async pulls)        */
  float* pubsmall[DPS_MAX_RANKS];        /* owner: published fp32 small prefix [n_small]         */
  uint32_t* flags[DPS_MAX_RANKS];        /* flag block [DPS_FLAG_WORDS]                          */
  /* Optional device-resident step counters and learning rate, so a captured CUDA graph can be replayed:
   * when non-NULL they override the `step` / `lr` arguments.  step_ctr[0] = sync round (bumped by
   * dps_wait_pull with bump=1), step_ctr[1] = this rank's async push count (bumped by dps_async_apply). */
  uint32_t* step_ctr;
  const float* lr_ptr;
  /* Optional NVLS window: the multicast address of the bf16 weight arena / fp32 small prefix.  One multimem.st through
   * it lands in EVERY worker's wbf16 / small (same offsets), so the pull stores each vector once instead of W times.
   * NULL = per-peer stores. */
  uint16_t* mc_wbf16;
  float* mc_small;
  /* Optional NVLS reduction of the sync push: when non-NULL, `stage` lives in the multicast-bound arena, laid out
   * [owner][E] on EVERY rank; SEND writes the source's fp16 tiles into its OWN replica (no peer stores) and the owner's
   * REDUCE reads mc_stage + owner*E + l with multimem.ld_reduce, which returns the sum over all W replicas.  fp16 wire
   * only, no extra sources (n_extra == 0). */
  uint16_t* mc_stage;
} DpsTables;

const char* dps_last_error(void);
int dps_tables_size(void);                 /* sizeof(DpsTables): bindings check their mirror of the struct against it */

/* ---- fabric (reference: the gRPC channel/server runtime, server.py:370-403, worker.py:199-231) ---- */
int dps_device_count(int* n);
int dps_arena_alloc(int device, uint64_t bytes, void** out_ptr);          /* cudaMalloc + zero          */
int dps_arena_free(int device, void* ptr);
int dps_ipc_export(void* ptr, unsigned char out_handle[64]);
int dps_ipc_open(int device, const unsigned char handle[64], void** out_ptr);
int dps_ipc_close(void* ptr);
int dps_enable_peer(int device, int peer);                                  /* 0 if (already) enabled     */

/* ---- PS data path ---- */
/* reference: worker.py:264-301 + server.py:232-288,126-169 (compress, push, decompress, mean, SGD) */
int dps_push_sync(const DpsTables* T, unsigned step, float lr, float momentum, float weight_decay,
                  int compress_fp16, int ctas, unsigned long long timeout_ns, void* stream);
/* The same kernel on the LOCAL shard range [seg_lo, seg_hi) (multiples of 8), phases: 1 = SEND only (never waits:
 * launched per layer group during backward), 2 = REDUCE only (waits for arrivals_per_step signals per live source
 * and round, then mean + SGD), 3 = both.  reference: the push happens once, after the whole backward (worker.py:376). */
int dps_push_segment(const DpsTables* T, unsigned step, float lr, float momentum, float weight_decay,
                     int compress_fp16, int ctas, unsigned long long timeout_ns, long long seg_lo, long long seg_hi,
                     int phases, unsigned arrivals_per_step, void* stream);
/* reference: server.py:213-230 + worker.py:233-262 (FetchParameters + load_state_dict) */
int dps_pull_bcast(const DpsTables* T, unsigned step, int ctas, void* stream);
int dps_wait_pull(const DpsTables* T, int rank, unsigned step, int ctas_per_owner, unsigned long long timeout_ns,
                  int bump, void* stream);
/* Segmented pull (layer groups in forward order) + per-group wait: overlaps the pull with the next forward pass.
 * seg_offsets: nseg+1 LOCAL shard offsets (multiples of 8).  dps_bump_step advances the device round counter. */
int dps_pull_segments(const DpsTables* T, const long long* seg_offsets, int nseg, int ctas, void* stream);
int dps_wait_segment(const DpsTables* T, int rank, int seg, unsigned step, int ctas_per_owner,
                     unsigned long long timeout_ns, void* stream);
int dps_bump_step(const DpsTables* T, void* stream);
/* reference: server.py:290-304,171-186 (async apply-on-arrival with staleness weighting) */
int dps_push_async(const DpsTables* T, int rank, unsigned step, int ctas, unsigned long long timeout_ns,
                   void* stream);
/* policy: 0 reference (version per accepted update, bound as given) | 1 bound widened by W-1 | 2 one version per
 * drained batch.  bump_step: 1 = this scan closes the rank's own push (same stream); 0 = a dedicated server stream
 * drives the apply and dps_pull_async(bump_step=1) closes the push instead. */
int dps_async_apply(const DpsTables* T, int owner, int producer_ctas, int staleness_bound, float alpha,
                    float floor_w, int max_batch, float lr, float momentum, float weight_decay, int ctas,
                    int policy, int bump_step, void* stream);
int dps_pull_async(const DpsTables* T, int rank, int ctas, int bump_step, void* stream);
int dps_publish(const DpsTables* T, int owner, void* stream);
/* *word = (*word & and_mask) | or_mask with device atomics (membership commands of the gRPC front door on flag
 * words that the reduce kernels also update; reference: the registry of server.py:190-211,306-318). */
int dps_flag_update(uint32_t* word, unsigned and_mask, unsigned or_mask, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPS_B200_H */
