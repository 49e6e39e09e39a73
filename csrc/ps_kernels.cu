// Parameter-server data path on NVLink peer memory: Push (sync), Pull, Async enqueue / scan / apply.
//
// One process per GPU; every GPU is a worker AND the owner of one shard of the parameter store.
// The flat parameter arena (forward-ordered, see store/shardmap.py) is sharded BLOCK-CYCLICALLY:
// global block k (BS elements) lives on owner k % W at local block k / W, so a pull streams the
// arena front-to-back from all owners at once.
//
// What each kernel replaces in the reference (src/parameter_server/server.py, src/workers/worker.py):
//   push_sync   : compress_gradients (worker.py:264-268) + pickle/gRPC push (worker.py:289-301) +
//                 decompress (server.py:232-237) + W-way mean (server.py:145-169) + SGD (server.py:126-143)
//                 -> one kernel: fp32 grad -> fp16 tiles stored straight into the owner's staging slot
//                 over NVLink, per-source arrival counters (release/acquire at system scope), then the
//                 owner reduces in registers and updates its fp32 master shard (+momentum, +weight decay).
//   pull_bcast  : pickle.dumps(params)/FetchParameters/load_state_dict (server.py:213-230, worker.py:233-262)
//                 -> owner casts its updated shard to bf16 and stores it into every worker's weight arena.
//   async_*     : the gRPC thread pool + param_lock + staleness logic (server.py:171-186, 290-304)
//                 -> a device-side arrival ring per owner (ticketed slots), a per-shard version counter,
//                 and a scan/apply pair that applies arrivals in ticket order with
//                 weight = max(floor, 1/(1+alpha*staleness)) or rejects them (staleness > bound).
//
// No NCCL and no host round trip on this path.  All cross-GPU flags use st.release.sys / ld.acquire.sys.
#include "common.cuh"
#include "../include/dps_b200.h"

thread_local char g_dps_err[512] = {0};

namespace dps {

constexpr int VEC = 8;  // elements per thread per iteration (16 B of fp16/bf16, 32 B of fp32)

// local shard offset -> arena offset.  BS is a power of two (check_tables): shift and mask, no 64-bit division in the
// copy loops.
__device__ __forceinline__ long long global_index(long long l, int owner, int W, int BS) {
  const int sh = __ffs(BS) - 1;
  const long long lb = l >> sh;
  return ((lb * W + owner) << sh) + (l & (long long)(BS - 1));
}

struct F8 {
  float v[8];
};
__device__ __forceinline__ F8 load_f32x8(const float* p) {
  F8 r;
  uint4 a = ld_v4(p), b = ld_v4(p + 4);
  r.v[0] = __uint_as_float(a.x); r.v[1] = __uint_as_float(a.y); r.v[2] = __uint_as_float(a.z); r.v[3] = __uint_as_float(a.w);
  r.v[4] = __uint_as_float(b.x); r.v[5] = __uint_as_float(b.y); r.v[6] = __uint_as_float(b.z); r.v[7] = __uint_as_float(b.w);
  return r;
}
__device__ __forceinline__ void store_f32x8(float* p, const F8& r) {
  st_v4(p, make_uint4(__float_as_uint(r.v[0]), __float_as_uint(r.v[1]), __float_as_uint(r.v[2]), __float_as_uint(r.v[3])));
  st_v4(p + 4, make_uint4(__float_as_uint(r.v[4]), __float_as_uint(r.v[5]), __float_as_uint(r.v[6]), __float_as_uint(r.v[7])));
}
__device__ __forceinline__ uint4 to_f16x8(const F8& r) {
  return make_uint4(pack_f16x2(r.v[0], r.v[1]), pack_f16x2(r.v[2], r.v[3]), pack_f16x2(r.v[4], r.v[5]),
                    pack_f16x2(r.v[6], r.v[7]));
}
__device__ __forceinline__ uint4 to_bf16x8(const F8& r) {
  return make_uint4(pack_bf16x2(r.v[0], r.v[1]), pack_bf16x2(r.v[2], r.v[3]), pack_bf16x2(r.v[4], r.v[5]),
                    pack_bf16x2(r.v[6], r.v[7]));
}
__device__ __forceinline__ void acc_f16x8(F8& a, uint4 h, float w) {
  float2 f;
  f = unpack_f16x2(h.x); a.v[0] += w * f.x; a.v[1] += w * f.y;
  f = unpack_f16x2(h.y); a.v[2] += w * f.x; a.v[3] += w * f.y;
  f = unpack_f16x2(h.z); a.v[4] += w * f.x; a.v[5] += w * f.y;
  f = unpack_f16x2(h.w); a.v[6] += w * f.x; a.v[7] += w * f.y;
}

__device__ __forceinline__ uint32_t owner_mask(int W) { return W >= 32 ? 0xffffffffu : ((1u << W) - 1u); }

// Spin until *flag >= target (acquire, system scope).  Returns false on timeout.
__device__ __forceinline__ bool wait_ge(const uint32_t* flag, uint32_t target, unsigned long long timeout_ns) {
  if ((int32_t)(ld_acquire_sys(flag) - target) >= 0) return true;
  const uint64_t t0 = globaltimer_ns();
  while ((int32_t)(ld_acquire_sys(flag) - target) < 0) {
    __nanosleep(64);
    if (timeout_ns && globaltimer_ns() - t0 > timeout_ns) return false;
  }
  return true;
}

// Remote sources (gRPC front door): once per round and owner, ONE thread decides which attached remote workers have a
// complete, not yet consumed gradient in their staging slot (arrival counter >= (consumed + 1) * signals per push), marks
// them in DPS_F_INCLUDE for the reduce that follows in the same stream, and records the consuming round.
__global__ void remote_gate_kernel(const __grid_constant__ DpsTables T, uint32_t step_arg, uint32_t arrivals_per_push) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const uint32_t step = T.step_ctr ? T.step_ctr[0] : step_arg;
  for (int o = T.rank_lo; o < T.rank_hi && o < T.world; ++o) {
    uint32_t* f = T.flags[o];
    const uint32_t alive = f[DPS_F_ALIVE];
    uint32_t mask = 0;
    for (int x = 0; x < T.n_extra; ++x) {
      const int s = T.world + x;
      if (!((alive >> s) & 1u)) continue;
      const uint32_t consumed = f[DPS_F_CONSUMED + x];
      if ((int32_t)(ld_acquire_sys(f + DPS_F_ARRIVE + s) - (consumed + 1) * arrivals_per_push) >= 0) {
        mask |= 1u << s;
        f[DPS_F_CONSUMED + x] = consumed + 1;
        f[DPS_F_CONSUMED_ROUND + x] = step;
      }
    }
    f[DPS_F_INCLUDE] = mask;
  }
}

// ------------------------------------------------------------------------------------------
// PUSH (sync): scatter fp16 gradient tiles to owners, barrier on per-source arrival counters,
// W-way mean + SGD(momentum, weight decay) on the owner's fp32 master shard.
// grid = G CTAs (G <= resident CTAs: owner-side CTAs spin, so the whole grid must be co-resident).
//
// The kernel works on the LOCAL shard range [seg_lo, seg_hi) and runs the phases selected by `phases`:
//   bit 0 (SEND)   : compress + peer stores + arrival signal.  Never waits, so it can be launched per layer group
//                    from the backward pass as soon as that group's weight gradients are complete (the arena is in
//                    forward order; layer groups are contiguous ranges of every shard).
//   bit 1 (REDUCE) : wait until every live source has signalled `arrivals_per_step` times this round, then mean + SGD.
// One launch with phases = 3 over [0, E) is the monolithic push; the overlapped step sends fc / layer4 / layer3 /
// layer2 on a side stream during backward and ends with SEND(first group) + REDUCE(everything).
// Membership: DPS_F_ALIVE is this owner's mask of sources it still waits for.  A source that misses a barrier
// (timeout) is dropped from the mask for good -- later rounds neither wait for it nor read its stale staging slot --
// and the sticky error bit tells the host watchdog to checkpoint and restart the job.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) push_sync_kernel(const __grid_constant__ DpsTables T, uint32_t step_arg,
                                                        float lr_arg, float momentum, float weight_decay,
                                                        int compress_fp16, unsigned long long timeout_ns,
                                                        long long seg_lo, long long seg_hi, int phases,
                                                        uint32_t arrivals_per_step) {
  const uint32_t step = T.step_ctr ? T.step_ctr[0] : step_arg;
  const float lr = T.lr_ptr ? *T.lr_ptr : lr_arg;
  const int W = T.world, BS = T.block_elems;
  const long long E = T.shard_elems;
  const int G = gridDim.x, b = blockIdx.x;
  const long long per = (((seg_hi - seg_lo) / VEC + G - 1) / G) * VEC;       // elements per CTA per shard
  const long long lo = seg_lo + (long long)b * per, hi = min(seg_hi, lo + per);
  __shared__ uint32_t s_present;

  // ---- phase A: every rank handled by this launch sends its slice of every shard to that shard's owner
  for (int r = T.rank_lo; r < T.rank_hi && (phases & 1); ++r) {
    float* grad = T.grad[r];
    for (int oo = 0; oo < W; ++oo) {
      const int o = (r + 1 + oo) % W;                        // rotate so owners are hit evenly
      // peer-mapped staging slot [src=r] on the owner -- or, with the NVLS reduction, slot [owner=o] of this rank's OWN
      // replica of the multicast-bound staging arena (local stores only; the owner sums the replicas in the switch)
      uint16_t* dst = T.mc_stage ? T.stage[r] + (long long)o * E : T.stage[o] + (long long)r * E;
      // four independent 32-byte loads per thread before the first store: the loop is bound by load latency otherwise
      const long long stride = (long long)blockDim.x * VEC;
      for (long long l0 = lo + threadIdx.x * VEC; l0 < hi; l0 += 4 * stride) {
        F8 v[4];
        long long g[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long long l = l0 + j * stride;
          if (l < hi) { g[j] = global_index(l, o, W, BS); v[j] = load_f32x8(grad + g[j]); }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long long l = l0 + j * stride;
          if (l >= hi) break;
          if (!compress_fp16) {
            // uncompressed wire (compress = 'none'): fp32 straight into the staging slot, which is allocated twice as
            // large for this mode (slot stride 2E halfs == E floats)
            store_f32x8(reinterpret_cast<float*>(T.stage[o]) + (long long)r * E + l, v[j]);
          } else {
            st_v4(dst + l, to_f16x8(v[j]));
          }
          F8 z = {};                                          // gradients are consumed: zero for next step
          store_f32x8(grad + g[j], z);
        }
      }
    }
    // one system-scope fence for all W destinations, then relaxed arrivals (a release per owner would cost a
    // MEMBAR.SYS each): every store above happens-before the fence through the CTA barrier
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      for (int o = 0; o < W; ++o) {
        red_relaxed_sys_add(T.flags[o] + DPS_F_ARRIVE + r, 1u);
        if (b == 0) st_relaxed_sys(T.flags[o] + DPS_F_HEARTBEAT + r, step + 1);
      }
    }
    __syncthreads();
  }

  // ---- phase B: owners wait for all live sources, reduce and update
  if (!(phases & 2)) return;
  const uint32_t target = (step + 1) * arrivals_per_step;
  for (int o = T.rank_lo; o < T.rank_hi; ++o) {
    if (threadIdx.x < 32) {
      // warp-parallel poll: lane s watches source s (W + extra <= 32), so a round costs ONE flag latency and a
      // missing source costs ONE timeout, not W of them in sequence.  Sources W.. are remote workers attached through
      // the gRPC front door (rpc/device_door.py): the owners never WAIT for them -- a CPU worker needs seconds per
      // step -- they are folded into whichever round is current once their gradient has arrived completely; that
      // decision is taken once per round by remote_gate_kernel (launched in front of this kernel) and read here from
      // DPS_F_INCLUDE, so all CTAs of the owner agree.
      const int s = threadIdx.x;
      uint32_t* f = T.flags[o];
      if (b == 0 && s == 0) st_relaxed_sys(f + DPS_F_ROUND, step);      // host-visible progress
      const uint32_t alive = f[DPS_F_ALIVE];
      const bool expected = s < W && ((alive >> s) & 1u);
      bool ok = false;
      if (expected) ok = wait_ge(f + DPS_F_ARRIVE + s, target, timeout_ns);
      if (s >= W && s < W + T.n_extra) ok = (f[DPS_F_INCLUDE] >> s) & 1u;
      const uint32_t present = __ballot_sync(0xffffffffu, ok);
      const uint32_t gone = __ballot_sync(0xffffffffu, expected && !ok);
      if (s == 0) {
        if (gone) {                                            // somebody we still counted on did not arrive
          atomicOr(f + DPS_F_ERROR, DPS_ERR_BARRIER_TIMEOUT);
          atomicOr(f + DPS_F_MISSING, gone);
          atomicAnd(f + DPS_F_ALIVE, ~gone);
        }
        s_present = present;
      }
    }
    __syncthreads();
    // device-side timing of the update itself (after the barrier wait): CTA 0's span is accumulated in DPS_F_UPDATE_NS
    // so the server metrics report a measured average update time (reference: update_times, server.py:121,128,140-143)
    const uint64_t t_upd0 = (b == 0 && threadIdx.x == 0) ? globaltimer_ns() : 0;
    const uint32_t present = s_present;
    const float inv = 1.0f / (float)max(1, __popc(present));
    float* master = T.master[o];
    float* mom = T.mom[o];
    for (long long l = lo + threadIdx.x * VEC; l < hi; l += blockDim.x * VEC) {
      F8 g = {};
      if (T.mc_stage) {
        // ONE load: the switch adds the W replicas (fp32 accumulation) and returns the sum.  A dropped source cannot be
        // left out of an in-switch sum; the sticky error bit raised above already makes the host restart the job.
        acc_f16x8(g, multimem_ld_reduce_f16x8(T.mc_stage + (long long)o * E + l), 1.0f);
      }
      for (int s = 0; s < W + T.n_extra && !T.mc_stage; ++s) {
        if (!((present >> s) & 1u)) continue;
        if (compress_fp16) {
          acc_f16x8(g, ld_v4(T.stage[o] + (long long)s * E + l), 1.0f);
        } else {
          F8 t = load_f32x8(reinterpret_cast<const float*>(T.stage[o]) + (long long)s * E + l);
#pragma unroll
          for (int i = 0; i < 8; ++i) g.v[i] += t.v[i];
        }
      }
      F8 p = load_f32x8(master + l);
      if (momentum != 0.0f) {
        F8 m = load_f32x8(mom + l);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float gi = g.v[i] * inv + weight_decay * p.v[i];
          m.v[i] = momentum * m.v[i] + gi;
          p.v[i] = p.v[i] - lr * m.v[i];
        }
        store_f32x8(mom + l, m);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float gi = g.v[i] * inv + weight_decay * p.v[i];
          p.v[i] = p.v[i] - lr * gi;                        // == reference p - lr*g when wd == 0
        }
      }
      store_f32x8(master + l, p);
    }
    __syncthreads();
    if (threadIdx.x == 0 && b == 0) {
      T.flags[o][DPS_F_VERSION] += 1;   // shard version (survives checkpoint resume)
      *reinterpret_cast<unsigned long long*>(T.flags[o] + DPS_F_UPDATE_NS) += globaltimer_ns() - t_upd0;
    }
  }
}

// NVLS form of the broadcast: ONE multimem.st per vector through the multicast window; the switch replicates it into
// every worker's arena (this GPU's included), so the owner's egress is E*2 bytes instead of (W-1)*E*2.
__device__ __forceinline__ void bcast_mc(const DpsTables& T, long long g, long long n_small, uint4 h, const F8& p) {
  multimem_st_v4(T.mc_wbf16 + g, h);
  if (g < n_small) {
    multimem_st_v4(T.mc_small + g, make_uint4(__float_as_uint(p.v[0]), __float_as_uint(p.v[1]), __float_as_uint(p.v[2]),
                                              __float_as_uint(p.v[3])));
    multimem_st_v4(T.mc_small + g + 4, make_uint4(__float_as_uint(p.v[4]), __float_as_uint(p.v[5]),
                                                  __float_as_uint(p.v[6]), __float_as_uint(p.v[7])));
  }
}

// ------------------------------------------------------------------------------------------
// PULL: owner casts its fp32 shard to bf16 and stores it into every worker's weight arena
// (plus an fp32 copy of the "small" head of the arena: BN affine parameters and fc bias).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pull_bcast_kernel(const __grid_constant__ DpsTables T, uint32_t step_arg) {
  const uint32_t step = T.step_ctr ? T.step_ctr[0] : step_arg;
  (void)step;
  const int W = T.world, BS = T.block_elems;
  const long long E = T.shard_elems, n_small = T.n_small;
  const int G = gridDim.x, b = blockIdx.x;
  const long long per = ((E / VEC + G - 1) / G) * VEC;
  const long long lo = (long long)b * per, hi = min(E, lo + per);
  for (int o = T.rank_lo; o < T.rank_hi; ++o) {
    const float* master = T.master[o];
    for (long long l = lo + threadIdx.x * VEC; l < hi; l += blockDim.x * VEC) {
      const long long g = global_index(l, o, W, BS);
      const F8 p = load_f32x8(master + l);
      const uint4 h = to_bf16x8(p);
      if (T.mc_wbf16) {
        bcast_mc(T, g, n_small, h, p);
        continue;
      }
      for (int ww = 0; ww < W; ++ww) {
        const int w = (o + ww) % W;                          // start with the local copy, then peers
        st_v4(T.wbf16[w] + g, h);
        if (g < n_small) store_f32x8(T.small[w] + g, p);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();                                  // one fence, then relaxed signals to all workers
      for (int w = 0; w < W; ++w) red_relaxed_sys_add(T.flags[w] + DPS_F_PULLCOUNT, 1u);
    }
  }
}

// Segmented PULL: the arena is in forward order, so the owner broadcasts it front to back in `nseg` layer groups
// and bumps a per-group delivery counter on every worker after each group.  The next forward pass waits per group
// (wait_seg_kernel in front of the group's first conv), so the pull overlaps the forward pass layer group by layer
// group instead of preceding it.  seg[] holds nseg+1 LOCAL element offsets (multiples of 8) into the shard.
struct SegBounds {
  long long off[DPS_MAX_SEGS + 1];
  int nseg;
};
__global__ void __launch_bounds__(256) pull_seg_kernel(const __grid_constant__ DpsTables T,
                                                       const __grid_constant__ SegBounds sb) {
  const int W = T.world, BS = T.block_elems;
  const long long n_small = T.n_small;
  const int G = gridDim.x, b = blockIdx.x;
  for (int sg = 0; sg < sb.nseg; ++sg) {
    const long long s0 = sb.off[sg], s1 = sb.off[sg + 1];
    const long long per = (((s1 - s0) / VEC + G - 1) / G) * VEC;
    const long long lo = s0 + (long long)b * per, hi = min(s1, lo + per);
    for (int o = T.rank_lo; o < T.rank_hi; ++o) {
      const float* master = T.master[o];
      const long long stride = (long long)blockDim.x * VEC;
      for (long long l0 = lo + threadIdx.x * VEC; l0 < hi; l0 += 4 * stride) {
        F8 p[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (l0 + j * stride < hi) p[j] = load_f32x8(master + l0 + j * stride);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long long l = l0 + j * stride;
          if (l >= hi) break;
          const long long g = global_index(l, o, W, BS);
          const uint4 h = to_bf16x8(p[j]);
          if (T.mc_wbf16) {
            bcast_mc(T, g, n_small, h, p[j]);
            continue;
          }
          for (int ww = 0; ww < W; ++ww) {
            const int w = (o + ww) % W;
            st_v4(T.wbf16[w] + g, h);
            if (g < n_small) store_f32x8(T.small[w] + g, p[j]);
          }
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();
        for (int w = 0; w < W; ++w) red_relaxed_sys_add(T.flags[w] + DPS_F_SEGCOUNT + sg, 1u);
      }
      __syncthreads();
    }
  }
}

// Deliveries expected so far are ACCUMULATED on the device (DPS_F_SEGEXPECT / DPS_F_PULLEXPECT += live owners x CTAs
// per round) instead of being derived from the round number, so owners dropped from DPS_F_ALIVE stop being waited for.
__global__ void wait_seg_kernel(const __grid_constant__ DpsTables T, int rank, int seg, uint32_t step_arg,
                                uint32_t ctas_per_owner, unsigned long long timeout_ns) {
  if (threadIdx.x == 0) {
    (void)step_arg;
    uint32_t* f = T.flags[rank];
    const uint32_t live = (uint32_t)__popc(f[DPS_F_ALIVE] & owner_mask(T.world));   // owners only: extra sources own no shard
    const uint32_t target = f[DPS_F_SEGEXPECT + seg] + ctas_per_owner * live;
    f[DPS_F_SEGEXPECT + seg] = target;
    if (!wait_ge(f + DPS_F_SEGCOUNT + seg, target, timeout_ns)) atomicOr(f + DPS_F_ERROR, DPS_ERR_PULL_TIMEOUT);
  }
}

// Last kernel of a round when the pull runs at the START of the step: advance the device-resident round counter.
__global__ void bump_step_kernel(const __grid_constant__ DpsTables T) {
  if (threadIdx.x == 0 && blockIdx.x == 0 && T.step_ctr) T.step_ctr[0] += 1;
}

// Worker side of the pull: block the stream until all W owners have delivered `step`.
__global__ void wait_pull_kernel(const __grid_constant__ DpsTables T, int rank, uint32_t step_arg,
                                 uint32_t ctas_per_owner, unsigned long long timeout_ns, int bump) {
  if (threadIdx.x == 0) {
    const uint32_t step = T.step_ctr ? T.step_ctr[0] : step_arg;
    uint32_t* f = T.flags[rank];
    const uint32_t target = f[DPS_F_PULLEXPECT] + ctas_per_owner * (uint32_t)__popc(f[DPS_F_ALIVE] & owner_mask(T.world));
    f[DPS_F_PULLEXPECT] = target;
    if (!wait_ge(f + DPS_F_PULLCOUNT, target, timeout_ns)) atomicOr(f + DPS_F_ERROR, DPS_ERR_PULL_TIMEOUT);
    if (bump && T.step_ctr) T.step_ctr[0] = step + 1;     // last kernel of the round: safe to advance
  }
}

// ------------------------------------------------------------------------------------------
// ASYNC: arrival ring.  grid = (G, W): blockIdx.y selects the owner.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) push_async_kernel(const __grid_constant__ DpsTables T, int r,
                                                         uint32_t step_arg, unsigned long long timeout_ns) {
  const uint32_t step = T.step_ctr ? T.step_ctr[1] : step_arg;
  const int W = T.world, BS = T.block_elems, R = T.ring_slots;
  const long long E = T.shard_elems;
  const int G = gridDim.x, b = blockIdx.x;
  const int o = (r + 1 + blockIdx.y) % W;
  const long long per = ((E / VEC + G - 1) / G) * VEC;
  const long long lo = (long long)b * per, hi = min(E, lo + per);
  uint32_t* lf = T.flags[r];                                  // my local flag block
  uint32_t* of = T.flags[o];                                  // owner's flag block (peer)
  __shared__ uint32_t s_ticket, s_drop;
  if (threadIdx.x == 0) {
    uint32_t ticket = 0, drop = 0;
    if (b == 0) {
      // WAIT-FREE admission: a push never blocks on another rank (a rank stuck in push could not serve its own
      // ring, and two such ranks would wait on each other).  Take a ticket only if the ring certainly has room:
      // at most W producers can pass this check concurrently, hence the margin of W slots.  Otherwise the push to
      // this shard is skipped for this step; the gradient slice is NOT zeroed, so it rides along with the next one.
      const uint32_t tail = ld_acquire_sys(of + DPS_F_RING_TAIL), head = ld_acquire_sys(of + DPS_F_RING_HEAD);
      if ((int32_t)(tail - head) > R - W) {
        drop = 1;
        atomicAdd(lf + DPS_F_DROPPED, 1u);
      } else {
        ticket = atom_add_acqrel_sys(of + DPS_F_RING_TAIL, 1u);
        uint32_t* hdr = of + DPS_F_SLOTS + (ticket % R) * 4;
        hdr[1] = (uint32_t)r;
        hdr[2] = lf[DPS_F_SEEN + o];                          // version of shard o this worker last pulled
      }
      // publish the decision to the other CTAs of this (rank, owner) pair
      lf[DPS_F_TICKET + 2 * o + 1] = ticket;
      lf[DPS_F_DROP + o] = drop;
      __threadfence();
      st_release_sys(lf + DPS_F_TICKET + 2 * o, step + 1);
    } else {
      while (ld_acquire_sys(lf + DPS_F_TICKET + 2 * o) != step + 1) __nanosleep(32);
      ticket = lf[DPS_F_TICKET + 2 * o + 1];
      drop = lf[DPS_F_DROP + o];
    }
    s_ticket = ticket;
    s_drop = drop;
  }
  __syncthreads();
  if (s_drop) return;
  const uint32_t ticket = s_ticket;
  const uint32_t slot = ticket % (uint32_t)R;
  uint16_t* dst = T.ring[o] + (long long)slot * E;
  float* grad = T.grad[r];
  const long long stride = (long long)blockDim.x * VEC;
  for (long long l0 = lo + threadIdx.x * VEC; l0 < hi; l0 += 4 * stride) {     // four loads in flight per thread
    F8 v[4];
    long long g[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (l0 + j * stride < hi) { g[j] = global_index(l0 + j * stride, o, W, BS); v[j] = load_f32x8(grad + g[j]); }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long l = l0 + j * stride;
      if (l >= hi) break;
      st_v4(dst + l, to_f16x8(v[j]));
      F8 z = {};
      store_f32x8(grad + g[j], z);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    red_release_sys_add(of + DPS_F_SLOTS + slot * 4, 1u);     // slot.done += 1 (complete at (use+1)*G)
    if (b == 0) st_release_sys(of + DPS_F_HEARTBEAT + r, step + 1);
  }
}

// Owner: examine the ring in ticket order, decide accept/reject + weight for each complete arrival.
// policy (staleness accounting, recorded in the metrics):
//   0 "reference"   : every accepted update bumps the shard version; reject staleness > bound (server.py:171-186)
//   1 "world"       : same, but the bound is widened by W - 1: with W workers that each pull once per step, W - 1 other
//                     updates land between a worker's pull and its push BY CONSTRUCTION, so the reference bound of 5
//                     silently discards ~40 % of the work at W = 8
//   2 "batch"       : all arrivals drained in one pass count as ONE version (they were concurrent)
// bump_step: 1 when the scan runs in the worker's own stream right after its push (the original schedule); 0 when a
// dedicated server stream drives it (the worker's pull then advances the push counter).
__global__ void async_scan_kernel(const __grid_constant__ DpsTables T, int o, uint32_t producer_ctas, int bound,
                                  float alpha, float floor_w, int max_batch, int policy, int bump_step) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  uint32_t* f = T.flags[o];
  const uint32_t R = (uint32_t)T.ring_slots;
  if (bump_step && T.step_ctr) T.step_ctr[1] += 1;        // this rank's push (same stream, before) is done
  if (policy == 1) bound += T.world - 1;
  uint32_t h = f[DPS_F_RING_HEAD], v = f[DPS_F_VERSION];
  const uint32_t v_pass = v;
  int n = 0, n_apply = 0;
  while (n < max_batch && n < DPS_MAX_PLAN) {
    const uint32_t slot = h % R;
    const uint32_t need = (h / R + 1) * producer_ctas;
    if ((int32_t)(ld_acquire_sys(f + DPS_F_SLOTS + slot * 4) - need) < 0) break;
    const uint32_t src = f[DPS_F_SLOTS + slot * 4 + 1];
    const uint32_t seen = f[DPS_F_SLOTS + slot * 4 + 2];
    const int s = (int)((policy == 2 ? v_pass : v) - seen);
    float w = 0.0f;
    if (s > bound) {
      f[DPS_F_STAT_REJECTED] += 1;
    } else {
      w = fmaxf(floor_w, 1.0f / (1.0f + alpha * (float)s));
      if (policy != 2 || n_apply == 0) v += 1;
      n_apply += 1;
      f[DPS_F_STAT_ACCEPTED] += 1;
    }
    const int sc = max(0, s);
    f[DPS_F_STAT_SUM] += (uint32_t)sc;
    f[DPS_F_STAT_MAX] = max(f[DPS_F_STAT_MAX], (uint32_t)sc);
    f[DPS_F_STAT_HIST + min(sc, DPS_HIST_BINS - 1)] += 1;
    if (src < 32) f[DPS_F_STAT_LAST + src] = (uint32_t)sc;
    f[DPS_F_PLAN + 3 + n] = slot;
    reinterpret_cast<float*>(f)[DPS_F_PLAN + 3 + DPS_MAX_PLAN + n] = w;
    ++n;
    ++h;
  }
  f[DPS_F_PLAN + 0] = (uint32_t)n;
  f[DPS_F_PLAN + 1] = v;
  f[DPS_F_PLAN + 2] = h;
  f[DPS_F_APPLY_DONE] = 0;
  if (n_apply > 0) st_release_sys(f + DPS_F_PUBSEQ, f[DPS_F_PUBSEQ] | 1u);   // odd: shard being rewritten
}

// Owner: apply the planned arrivals sequentially per element (momentum order preserved), publish bf16.
__global__ void __launch_bounds__(256) async_apply_kernel(const __grid_constant__ DpsTables T, int o, float lr_arg,
                                                          float momentum, float weight_decay) {
  const float lr = T.lr_ptr ? *T.lr_ptr : lr_arg;
  uint32_t* f = T.flags[o];
  const int n = (int)f[DPS_F_PLAN + 0];
  if (n == 0) return;
  const int W = T.world, BS = T.block_elems;
  const long long E = T.shard_elems, n_small = T.n_small;
  const int G = gridDim.x, b = blockIdx.x;
  const long long per = ((E / VEC + G - 1) / G) * VEC;
  const long long lo = (long long)b * per, hi = min(E, lo + per);
  __shared__ uint32_t s_slot[DPS_MAX_PLAN];
  __shared__ float s_w[DPS_MAX_PLAN];
  if (threadIdx.x < n) {
    s_slot[threadIdx.x] = f[DPS_F_PLAN + 3 + threadIdx.x];
    s_w[threadIdx.x] = reinterpret_cast<float*>(f)[DPS_F_PLAN + 3 + DPS_MAX_PLAN + threadIdx.x];
  }
  __syncthreads();
  bool any = false;
  for (int k = 0; k < n; ++k) any |= (s_w[k] > 0.0f);
  const uint64_t t_upd0 = (b == 0 && threadIdx.x == 0) ? globaltimer_ns() : 0;
  float* master = T.master[o];
  float* mom = T.mom[o];
  if (any) {
    for (long long l = lo + threadIdx.x * VEC; l < hi; l += blockDim.x * VEC) {
      F8 p = load_f32x8(master + l);
      F8 m = {};
      if (momentum != 0.0f) m = load_f32x8(mom + l);
      for (int k = 0; k < n; ++k) {
        const float w = s_w[k];
        if (w <= 0.0f) continue;
        F8 g = {};
        acc_f16x8(g, ld_v4(T.ring[o] + (long long)s_slot[k] * E + l), w);   // staleness-weighted gradient
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float gi = g.v[i] + weight_decay * p.v[i];
          if (momentum != 0.0f) { m.v[i] = momentum * m.v[i] + gi; gi = m.v[i]; }
          p.v[i] = p.v[i] - lr * gi;
        }
      }
      store_f32x8(master + l, p);
      if (momentum != 0.0f) store_f32x8(mom + l, m);
      const long long g = global_index(l, o, W, BS);
      st_v4(T.pub[o] + l, to_bf16x8(p));
      if (g < n_small) store_f32x8(T.pubsmall[o] + g, p);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (b == 0 && any) *reinterpret_cast<unsigned long long*>(f + DPS_F_UPDATE_NS) += globaltimer_ns() - t_upd0;
    __threadfence();
    const uint32_t done = atomicAdd(f + DPS_F_APPLY_DONE, 1u) + 1;
    if (done == (uint32_t)G) {                                 // last CTA publishes
      __threadfence();
      f[DPS_F_VERSION] = f[DPS_F_PLAN + 1];
      if (any) st_release_sys(f + DPS_F_PUBSEQ, (f[DPS_F_PUBSEQ] | 1u) + 1u);   // even: consistent again
      st_release_sys(f + DPS_F_RING_HEAD, f[DPS_F_PLAN + 2]);  // frees the consumed slots
      f[DPS_F_PLAN + 0] = 0;
    }
  }
}

// Worker: gather every owner's published bf16 shard (peer reads) under a per-chunk seq-lock and record
// the version seen (the `local_step` of the next push, reference worker.py:245,299).  grid = (G, W).
__global__ void __launch_bounds__(256) pull_async_kernel(const __grid_constant__ DpsTables T, int r, int bump_step) {
  // bump_step: this pull follows the rank's push in the same stream and closes its step (dedicated-server schedule)
  if (bump_step && T.step_ctr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) T.step_ctr[1] += 1;
  const int W = T.world, BS = T.block_elems;
  const long long E = T.shard_elems, n_small = T.n_small;
  const int G = gridDim.x, b = blockIdx.x;
  const int o = (r + blockIdx.y) % W;
  const long long per = ((E / VEC + G - 1) / G) * VEC;
  const long long lo = (long long)b * per, hi = min(E, lo + per);
  const uint32_t* of = T.flags[o];
  __shared__ uint32_t s_seq, s_ok, s_ver;
  for (int attempt = 0; attempt < 4; ++attempt) {
    if (threadIdx.x == 0) {
      uint32_t s0 = ld_acquire_sys(of + DPS_F_PUBSEQ);
      int spins = 0;
      while ((s0 & 1u) && spins++ < (1 << 16)) { __nanosleep(64); s0 = ld_acquire_sys(of + DPS_F_PUBSEQ); }
      s_seq = s0;
      // the version that goes with this sequence value: read AFTER the sequence, before the copy; it is recorded as
      // "seen" only if the sequence is unchanged after the copy, so a worker never reports a newer version than the
      // data it holds (that would under-state the staleness of its next push)
      s_ver = ld_acquire_sys(of + DPS_F_VERSION);
    }
    __syncthreads();
    // peer LOADS over NVLink have ~2 us of latency: eight 16-byte loads per thread in flight before the first store
    // (one at a time, the 22 MB gather cost ~90 us at 4 GPUs -- most of the async step's gap to the sync step)
    const long long stride = (long long)blockDim.x * VEC;
    for (long long l0 = lo + threadIdx.x * VEC; l0 < hi; l0 += 8 * stride) {
      uint4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (l0 + j * stride < hi) v[j] = ld_v4(T.pub[o] + l0 + j * stride);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const long long l = l0 + j * stride;
        if (l >= hi) break;
        const long long g = global_index(l, o, W, BS);
        st_v4(T.wbf16[r] + g, v[j]);
        if (g < n_small) store_f32x8(T.small[r] + g, load_f32x8(T.pubsmall[o] + g));
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      // an odd sequence (the owner was still rewriting when the bounded spin gave up) is a failed attempt even if it
      // has not changed since
      s_ok = ((s_seq & 1u) == 0u && ld_acquire_sys(of + DPS_F_PUBSEQ) == s_seq) ? 1u : 0u;
      if (s_ok && b == 0) T.flags[r][DPS_F_SEEN + o] = s_ver;
      if (!s_ok && attempt == 3) {
        atomicAdd(T.flags[r] + DPS_F_TORN_PULLS, 1u);
        // keep the OLDEST version any attempt saw: over-stating staleness is safe, under-stating it is not
        if (b == 0) atomicMin(T.flags[r] + DPS_F_SEEN + o, s_ver);
      }
    }
    __syncthreads();
    if (s_ok) break;
  }
}

// Initial publish (and re-publish after a checkpoint load): master -> pub bf16 (+small fp32), no version bump.
__global__ void __launch_bounds__(256) publish_kernel(const __grid_constant__ DpsTables T, int o) {
  const int W = T.world, BS = T.block_elems;
  const long long E = T.shard_elems, n_small = T.n_small;
  for (long long l = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * VEC; l < E;
       l += (long long)gridDim.x * blockDim.x * VEC) {
    const F8 p = load_f32x8(T.master[o] + l);
    const long long g = global_index(l, o, W, BS);
    st_v4(T.pub[o] + l, to_bf16x8(p));
    if (g < n_small) store_f32x8(T.pubsmall[o] + g, p);
  }
}

}  // namespace dps

// ==========================================================================================
// C ABI
// ==========================================================================================
using namespace dps;

DPS_API const char* dps_last_error(void) { return g_dps_err; }
DPS_API int dps_tables_size(void) { return (int)sizeof(DpsTables); }   // checked against the ctypes mirror at load

static int check_tables(const DpsTables* T) {
  if (!T) DPS_FAIL(1, "null tables");
  if (T->world < 1 || T->world > DPS_MAX_RANKS) DPS_FAIL(1, "world %d out of range", T->world);
  if (T->n_extra < 0 || T->world + T->n_extra > DPS_MAX_RANKS) DPS_FAIL(1, "world + extra sources exceeds %d", DPS_MAX_RANKS);
  if (T->rank_lo < 0 || T->rank_hi > T->world + T->n_extra || T->rank_lo >= T->rank_hi) DPS_FAIL(1, "bad rank range");
  if (T->block_elems <= 0 || T->block_elems % 8 != 0 || (T->block_elems & (T->block_elems - 1)))
    DPS_FAIL(1, "block_elems must be a power of two >= 8");
  if (T->shard_elems <= 0 || T->shard_elems % T->block_elems != 0) DPS_FAIL(1, "shard_elems must be a multiple of block_elems");
  if (T->mc_stage && T->n_extra) DPS_FAIL(1, "the NVLS reduction sums every replica: no extra (remote) sources");
  return 0;
}

DPS_API int dps_push_sync(const DpsTables* T, unsigned step, float lr, float momentum, float weight_decay,
                          int compress_fp16, int ctas, unsigned long long timeout_ns, void* stream) {
  if (int e = check_tables(T)) return e;
  if (T->mc_stage && !compress_fp16) DPS_FAIL(1, "the NVLS reduction needs the fp16 wire format");
  if (T->n_extra > 0 && T->rank_lo < T->world)
    remote_gate_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T, step, (uint32_t)ctas);
  push_sync_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(*T, step, lr, momentum, weight_decay, compress_fp16,
                                                          timeout_ns, 0, T->shard_elems, 3, (uint32_t)ctas);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_push_segment(const DpsTables* T, unsigned step, float lr, float momentum, float weight_decay,
                             int compress_fp16, int ctas, unsigned long long timeout_ns, long long seg_lo,
                             long long seg_hi, int phases, unsigned arrivals_per_step, void* stream) {
  if (int e = check_tables(T)) return e;
  if (seg_lo < 0 || seg_hi > T->shard_elems || seg_lo > seg_hi || seg_lo % 8 || seg_hi % 8)
    DPS_FAIL(1, "push_segment: bad range [%lld, %lld)", seg_lo, seg_hi);
  if (!(phases & 3)) DPS_FAIL(1, "push_segment: phases must select SEND (1) and/or REDUCE (2)");
  if (T->mc_stage && !compress_fp16) DPS_FAIL(1, "the NVLS reduction needs the fp16 wire format");
  if (ctas < 1) DPS_FAIL(1, "push_segment: ctas must be positive");
  if ((phases & 2) && T->n_extra > 0 && T->rank_lo < T->world)
    remote_gate_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T, step, arrivals_per_step);
  push_sync_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(*T, step, lr, momentum, weight_decay, compress_fp16,
                                                          timeout_ns, seg_lo, seg_hi, phases, arrivals_per_step);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_pull_bcast(const DpsTables* T, unsigned step, int ctas, void* stream) {
  if (int e = check_tables(T)) return e;
  pull_bcast_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(*T, step);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_wait_pull(const DpsTables* T, int rank, unsigned step, int ctas_per_owner,
                          unsigned long long timeout_ns, int bump, void* stream) {
  if (int e = check_tables(T)) return e;
  wait_pull_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T, rank, step, (uint32_t)ctas_per_owner, timeout_ns, bump);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_pull_segments(const DpsTables* T, const long long* seg_offsets, int nseg, int ctas, void* stream) {
  if (int e = check_tables(T)) return e;
  if (nseg < 1 || nseg > DPS_MAX_SEGS) DPS_FAIL(1, "nseg out of range");
  SegBounds sb;
  sb.nseg = nseg;
  for (int i = 0; i <= nseg; ++i) {
    sb.off[i] = seg_offsets[i];
    if (sb.off[i] % 8 || sb.off[i] < 0 || sb.off[i] > T->shard_elems || (i && sb.off[i] < sb.off[i - 1]))
      DPS_FAIL(1, "bad segment offset %lld", (long long)sb.off[i]);
  }
  pull_seg_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(*T, sb);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_wait_segment(const DpsTables* T, int rank, int seg, unsigned step, int ctas_per_owner,
                             unsigned long long timeout_ns, void* stream) {
  if (int e = check_tables(T)) return e;
  wait_seg_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T, rank, seg, step, (uint32_t)ctas_per_owner, timeout_ns);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_bump_step(const DpsTables* T, void* stream) {
  if (int e = check_tables(T)) return e;
  bump_step_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_push_async(const DpsTables* T, int rank, unsigned step, int ctas, unsigned long long timeout_ns,
                           void* stream) {
  if (int e = check_tables(T)) return e;
  if (T->ring_slots < 1 || T->ring_slots > DPS_MAX_SLOTS) DPS_FAIL(1, "ring_slots out of range");
  dim3 grid(ctas, T->world);
  push_async_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*T, rank, step, timeout_ns);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_async_apply(const DpsTables* T, int owner, int producer_ctas, int staleness_bound, float alpha,
                            float floor_w, int max_batch, float lr, float momentum, float weight_decay, int ctas,
                            int policy, int bump_step, void* stream) {
  if (int e = check_tables(T)) return e;
  if (policy < 0 || policy > 2) DPS_FAIL(1, "async_apply: unknown staleness policy %d", policy);
  async_scan_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*T, owner, (uint32_t)producer_ctas, staleness_bound, alpha,
                                                         floor_w, max_batch, policy, bump_step);
  DPS_CHECK(cudaGetLastError());
  async_apply_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(*T, owner, lr, momentum, weight_decay);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_pull_async(const DpsTables* T, int rank, int ctas, int bump_step, void* stream) {
  if (int e = check_tables(T)) return e;
  dim3 grid(ctas, T->world);
  pull_async_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*T, rank, bump_step);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

// Host-side membership commands on a (possibly peer-mapped) flag word: *word = (*word & and_mask) | or_mask, atomically
// with respect to the kernels that update the same word (the reduce phase drops dead sources with atomicAnd).
namespace dps {
__global__ void flag_update_kernel(uint32_t* word, uint32_t and_mask, uint32_t or_mask) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    atomicAnd(word, and_mask);
    if (or_mask) atomicOr(word, or_mask);
    __threadfence_system();
  }
}
}  // namespace dps

DPS_API int dps_flag_update(uint32_t* word, unsigned and_mask, unsigned or_mask, void* stream) {
  if (!word) DPS_FAIL(1, "flag_update: null word");
  flag_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(word, and_mask, or_mask);
  DPS_CHECK(cudaGetLastError());
  return 0;
}

DPS_API int dps_publish(const DpsTables* T, int owner, void* stream) {
  if (int e = check_tables(T)) return e;
  publish_kernel<<<148, 256, 0, (cudaStream_t)stream>>>(*T, owner);
  DPS_CHECK(cudaGetLastError());
  return 0;
}
