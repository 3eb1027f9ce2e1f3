// Small-vector sum all-reduce over NVLink / NVSwitch peer memory, fused with the CG update.
//
// The multi-GPU update exchanges only short fp64 vectors (gradient d+1, each Fisher-vector
// product d, Gram (n+4)(n+5)+1, a few scalars; <= 107 KB), ~16 times per update, all latency bound.
// Instead of one NCCL launch per exchange, every rank keeps ONE symmetric block (cudaMalloc, exported with
// cudaIpc) that its peers map:
//     [ flags: MAXW u64, written by the peers ][ slot 0: max_count fp64 ][ slot 1: max_count fp64 ]
// An exchange at epoch e (a host counter all ranks advance in lock step) is ONE single-CTA kernel per rank:
//   1. the local contribution sits in slot e&1 (copied in, or written there by the producing kernel);
//   2. one thread per peer stores e into that peer's flags[self] (release, system scope);
//   3. one thread per peer spins on the LOCAL flags[peer] >= e (acquire, system scope; bounded by a timeout
//      that raises an error word instead of hanging the GPU);
//   4. every thread pulls its elements of all W slots over NVLink and adds them in RANK ORDER, so all ranks
//      hold bit-identical sums (the replicated CG state never diverges);
//   5. (fused variant) the CG update of cg_solve.py:10-20 runs on those sums in the same kernel.
// Slot e&1 is rewritten at epoch e+2, which a rank reaches only after every peer raised e+1, i.e. after every
// peer finished reading epoch e: two slots suffice and flags never need resetting (monotonic epochs).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace peer {

constexpr int MAXW = 8;
constexpr int THREADS = 1024;
constexpr int CG_E = 8;
constexpr size_t FLAG_BYTES = 256;   // MAXW u64 flags, padded

struct View {
  double* data[MAXW];                 // slot base of every rank (own entry = local pointer)
  unsigned long long* flags[MAXW];    // flag array of every rank
  int rank, world;
  long long max_count;
  int* err;                           // local: 1 = a wait timed out
  unsigned long long timeout_ns;
  unsigned long long* stamps;         // measurement aid (MJRL_B200_PEER_STAMPS=1), else null: [STAMP_EPOCHS][4] globaltimer ns
};
constexpr int STAMP_EPOCHS = 4096;
__device__ __forceinline__ unsigned long long now_ns();
// stamp i of this epoch's exchange kernel: 0 entry, 1 every peer's flag seen, 2 sums pulled (+ first dot product), 3 end
__device__ __forceinline__ void stamp(const View& pv, unsigned long long epoch, int i) {
  if (pv.stamps != nullptr && threadIdx.x == 0) pv.stamps[(epoch % STAMP_EPOCHS) * 4 + i] = now_ns();
}

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_peer(const double* p) {   // never served from a stale L1 line
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// steps 2-3; all threads of the CTA call it.  On return every peer's slot (epoch & 1) holds its
// contribution of this epoch (or *err is set).
__device__ __forceinline__ void handshake(const View& pv, unsigned long long epoch) {
  stamp(pv, epoch, 0);
  __threadfence_system();
  __syncthreads();
  const int t = threadIdx.x;
  if (t < pv.world && t != pv.rank) {
    st_release_sys(pv.flags[t] + pv.rank, epoch);
    const unsigned long long* mine = pv.flags[pv.rank] + t;
    const unsigned long long t0 = now_ns();
    while (ld_acquire_sys(mine) < epoch) {
      if (now_ns() - t0 > pv.timeout_ns) { *pv.err = 1; break; }
      __nanosleep(40);
    }
  }
  __syncthreads();
  __threadfence_system();
  stamp(pv, epoch, 1);
}

__device__ __forceinline__ double sum_ranks(const View& pv, long long off) {
  double v[MAXW];
#pragma unroll
  for (int r = 0; r < MAXW; ++r)
    if (r < pv.world) v[r] = ld_peer(pv.data[r] + off);
  double s = 0.0;
#pragma unroll
  for (int r = 0; r < MAXW; ++r)
    if (r < pv.world) s += v[r];
  return s;
}

// Pull with every load in flight at once.  A thread that sums W peer values per element in registers stalls on the first
// use, one NVLink round trip (~2 us) per element it owns: 13 us for d = 5,126 on 2 GPUs (measured, mjrl_peer_stamps).
// Instead the slots are staged in shared memory by 16-byte cp.async.cg (no registers, bypasses L1, so no stale line
// of an earlier epoch can be read; peer memory is not cached in the local L2): all copies of a chunk are issued
// before the first wait, then the sums are formed from shared memory in RANK ORDER as before.
constexpr long long BUF_ELEMS = 24576;           // 192 KB: [world][chunk] doubles
__device__ __forceinline__ long long stage_chunk_elems(int world) { return (BUF_ELEMS / world) & ~1023LL; }
// stage elements [e0, e0 + ce) of every rank's slot; returns after a CTA barrier (data visible to all threads)
__device__ __forceinline__ void stage_chunk(const View& pv, long long so, long long e0, long long ce, long long chunk,
                                            double* sbuf) {
  const long long c2 = (ce + 1) >> 1;            // 16-byte pieces per rank (may read one double past ce: slack exists)
  const long long total = (long long)pv.world * c2;
  for (long long c = threadIdx.x; c < total; c += THREADS) {
    const int r = (int)(c / c2);
    const long long j = c - (long long)r * c2;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                 :: "r"((uint32_t)__cvta_generic_to_shared(sbuf + r * chunk + 2 * j)), "l"(pv.data[r] + so + e0 + 2 * j) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
}
__device__ __forceinline__ double staged_sum(const View& pv, const double* sbuf, long long chunk, long long j) {
  double s = 0.0;
#pragma unroll
  for (int r = 0; r < MAXW; ++r)
    if (r < pv.world) s += sbuf[r * chunk + j];
  return s;
}

// dst[i] = sum_r slot_r[i]; src (may be NULL = already in the local slot) is copied into the slot first
__global__ void __launch_bounds__(THREADS)
allreduce_kernel(View pv, const double* __restrict__ src, double* __restrict__ dst, long long count,
                 unsigned long long epoch) {
  const long long so = (long long)(epoch & 1ull) * pv.max_count;
  double* mine = pv.data[pv.rank] + so;
  if (src != nullptr && src != mine)
    for (long long i = threadIdx.x; i < count; i += THREADS) mine[i] = src[i];
  handshake(pv, epoch);
  extern __shared__ __align__(16) double sbuf[];
  const long long chunk = stage_chunk_elems(pv.world);
  for (long long e0 = 0; e0 < count; e0 += chunk) {
    const long long ce = count - e0 < chunk ? count - e0 : chunk;
    stage_chunk(pv, so, e0, ce, chunk, sbuf);
    for (long long j = threadIdx.x; j < ce; j += THREADS) dst[e0 + j] = staged_sum(pv, sbuf, chunk, j);
    __syncthreads();
  }
  stamp(pv, epoch, 2);
  stamp(pv, epoch, 3);
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double cta_sum_bcast(double v, double* red) {
  v = warp_sum_d(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  double r = (l < nw) ? red[l] : 0.0;
  return warp_sum_d(r);
}

// The CG update (cg_solve.py:10-20, z = z_sum/count + damping*p, npg_cg.py:74) of csrc/policy.cu's
// cg_update_kernel with the all-reduce of z_sum folded in: same arithmetic and order after the sum.
// The local product sum must already be in slot epoch&1 (mjrl_policy_fvp* writes it there).
template <bool REGS>
__global__ void __launch_bounds__(THREADS)
cg_update_kernel(View pv, mjrl_cg_state st, long long d, double inv_count, double damping, double tol,
                 unsigned long long epoch) {
  __shared__ double red[32];
  const long long so = (long long)(epoch & 1ull) * pv.max_count;
  handshake(pv, epoch);          // also when converged: the peers count on this epoch's flag
  if (st.flags[0]) return;
  const double rdotr = st.scal[0];
  double pz = 0.0;
  double rp[REGS ? CG_E : 1], rz[REGS ? CG_E : 1], rx[REGS ? CG_E : 1], rr[REGS ? CG_E : 1];
  extern __shared__ __align__(16) double sbuf[];
  const long long chunk = stage_chunk_elems(pv.world);     // a multiple of THREADS: element i = tid + k THREADS keeps its k
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) rz[k] = 0.0;
    for (long long e0 = 0; e0 < d; e0 += chunk) {
      const long long ce = d - e0 < chunk ? d - e0 : chunk;
      stage_chunk(pv, so, e0, ce, chunk, sbuf);
#pragma unroll
      for (int k = 0; k < CG_E; ++k) {
        const long long j = threadIdx.x + (long long)k * THREADS - e0;
        if (j >= 0 && j < ce) rz[k] = staged_sum(pv, sbuf, chunk, j);
      }
      __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      const long long i = threadIdx.x + (long long)k * THREADS;
      const bool in = i < d;
      rp[k] = in ? st.p[i] : 0.0;
      rx[k] = in ? st.x[i] : 0.0; rr[k] = in ? st.r[i] : 0.0;
    }
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      rz[k] = fma(rz[k], inv_count, damping * rp[k]);
      pz = fma(rp[k], rz[k], pz);
    }
  } else {
    for (long long e0 = 0; e0 < d; e0 += chunk) {
      const long long ce = d - e0 < chunk ? d - e0 : chunk;
      stage_chunk(pv, so, e0, ce, chunk, sbuf);
      for (long long j = threadIdx.x; j < ce; j += THREADS) {
        const long long i = e0 + j;
        const double p = st.p[i];
        const double z = fma(staged_sum(pv, sbuf, chunk, j), inv_count, damping * p);
        st.z[i] = z;
        pz = fma(p, z, pz);
      }
      __syncthreads();
    }
  }
  pz = cta_sum_bcast(pz, red);
  stamp(pv, epoch, 2);
  const double v = rdotr / pz;
  double nrr = 0.0;
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      rx[k] = fma(v, rp[k], rx[k]);
      rr[k] = fma(-v, rz[k], rr[k]);
      nrr = fma(rr[k], rr[k], nrr);
    }
  } else {
    for (long long i = threadIdx.x; i < d; i += THREADS) {
      st.x[i] = fma(v, st.p[i], st.x[i]);
      const double r = fma(-v, st.z[i], st.r[i]);
      st.r[i] = r;
      nrr = fma(r, r, nrr);
    }
  }
  nrr = cta_sum_bcast(nrr, red);
  const double mu = nrr / rdotr;
  if (REGS) {
#pragma unroll
    for (int k = 0; k < CG_E; ++k) {
      const long long i = threadIdx.x + (long long)k * THREADS;
      if (i < d) {
        const double p = fma(mu, rp[k], rr[k]);
        st.z[i] = rz[k]; st.x[i] = rx[k]; st.r[i] = rr[k];
        st.p[i] = p; st.p32[i] = (float)p;
      }
    }
  } else {
    for (long long i = threadIdx.x; i < d; i += THREADS) {
      const double p = fma(mu, st.p[i], st.r[i]);
      st.p[i] = p; st.p32[i] = (float)p;
    }
  }
  if (threadIdx.x == 0) {
    st.scal[0] = nrr; st.scal[1] = pz; st.scal[2] = v; st.scal[3] = mu;
    st.flags[1] += 1;
    if (nrr < tol) st.flags[0] = 1;
  }
  stamp(pv, epoch, 3);
}

}  // namespace peer

struct mjrl_peer {
  int device, rank, world;
  long long max_count;
  void* base;                          // own block
  void* bases[peer::MAXW];             // mapped blocks of every rank (own entry = base)
  bool opened[peer::MAXW];             // mapped through cudaIpcOpenMemHandle (to be closed)
  bool connected;
  int* err;
  unsigned long long timeout_ns;
  unsigned long long* stamps;          // device [STAMP_EPOCHS][4] when MJRL_B200_PEER_STAMPS=1 at creation
  peer::View view;
};

static void fill_view(mjrl_peer* p) {
  peer::View& v = p->view;
  for (int r = 0; r < peer::MAXW; ++r) {
    char* b = (char*)(r < p->world ? p->bases[r] : nullptr);
    v.flags[r] = (unsigned long long*)b;
    v.data[r] = b ? (double*)(b + peer::FLAG_BYTES) : nullptr;
  }
  v.rank = p->rank; v.world = p->world; v.max_count = p->max_count; v.err = p->err;
  v.timeout_ns = p->timeout_ns;
  v.stamps = p->stamps;
}

// the exchange kernels stage the slots in 192 KB of dynamic shared memory
static int peer_kernel_attrs() {
  static bool done = false;
  if (done) return 0;
  const int bytes = (int)(peer::BUF_ELEMS * sizeof(double));
  cudaError_t e = cudaFuncSetAttribute(peer::allreduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(peer::cg_update_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(peer::cg_update_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) { mjrl_set_error("cudaFuncSetAttribute(peer kernels): %s", cudaGetErrorString(e)); return (int)e; }
  done = true;
  return 0;
}

extern "C" {

int mjrl_peer_create(mjrl_ctx* ctx, int32_t rank, int32_t world, int64_t max_count, mjrl_peer** out,
                     void* handle_out) {
  MJRL_REQUIRE(ctx && out && world >= 1 && world <= peer::MAXW && rank >= 0 && rank < world && max_count > 0,
               MJRL_E_ARG, "peer_create: bad argument (world <= %d)", peer::MAXW);
  cudaSetDevice(ctx->device);
  mjrl_peer* p = new mjrl_peer();
  memset(p, 0, sizeof(*p));
  max_count = (max_count + 1) & ~(int64_t)1;      // 16-byte aligned second slot (the exchange kernels copy 16-byte pieces)
  p->device = ctx->device; p->rank = rank; p->world = world; p->max_count = max_count;
  p->timeout_ns = 20ull * 1000ull * 1000ull * 1000ull;
  const size_t bytes = peer::FLAG_BYTES + 2 * (size_t)max_count * sizeof(double) + 16;   // + slack for the last 16-byte piece
  cudaError_t e = cudaMalloc(&p->base, bytes);
  if (e == cudaSuccess) e = cudaMemset(p->base, 0, bytes);
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->err, sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(p->err, 0, sizeof(int));
  if (e == cudaSuccess && getenv("MJRL_B200_PEER_STAMPS") && getenv("MJRL_B200_PEER_STAMPS")[0] == '1') {
    e = cudaMalloc((void**)&p->stamps, sizeof(unsigned long long) * 4 * peer::STAMP_EPOCHS);
    if (e == cudaSuccess) e = cudaMemset(p->stamps, 0, sizeof(unsigned long long) * 4 * peer::STAMP_EPOCHS);
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess && handle_out) {
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, p->base);
    if (e == cudaSuccess) memcpy(handle_out, &h, sizeof(h));
  }
  if (e != cudaSuccess) {
    mjrl_set_error("peer_create: %s", cudaGetErrorString(e));
    if (p->base) cudaFree(p->base);
    if (p->err) cudaFree(p->err);
    delete p;
    return (int)e;
  }
  p->bases[rank] = p->base;
  *out = p;
  return 0;
}

int mjrl_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

int mjrl_peer_connect(mjrl_peer* p, const void* handles) {
  MJRL_REQUIRE(p && handles && !p->connected, MJRL_E_ARG, "peer_connect: bad argument");
  cudaSetDevice(p->device);
  const char* h = (const char*)handles;
  for (int r = 0; r < p->world; ++r) {
    if (r == p->rank) continue;
    cudaIpcMemHandle_t hd;
    memcpy(&hd, h + (size_t)r * sizeof(hd), sizeof(hd));
    cudaError_t e = cudaIpcOpenMemHandle(&p->bases[r], hd, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      mjrl_set_error("peer_connect: cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
      cudaGetLastError();
      return (int)e;
    }
    p->opened[r] = true;
  }
  p->connected = true;
  fill_view(p);
  return 0;
}

int mjrl_peer_connect_ptrs(mjrl_peer* p, void* const* bases) {
  MJRL_REQUIRE(p && bases && !p->connected, MJRL_E_ARG, "peer_connect_ptrs: bad argument");
  for (int r = 0; r < p->world; ++r)
    if (r != p->rank) p->bases[r] = bases[r];
  p->connected = true;
  fill_view(p);
  return 0;
}

void* mjrl_peer_base(mjrl_peer* p) { return p ? p->base : nullptr; }

double* mjrl_peer_slot(mjrl_peer* p, int64_t epoch) {
  if (!p) return nullptr;
  return (double*)((char*)p->base + peer::FLAG_BYTES) + (size_t)(epoch & 1) * p->max_count;
}

int mjrl_peer_set_timeout(mjrl_peer* p, double seconds) {
  MJRL_REQUIRE(p && seconds > 0, MJRL_E_ARG, "peer_set_timeout: bad argument");
  p->timeout_ns = (unsigned long long)(seconds * 1e9);
  p->view.timeout_ns = p->timeout_ns;
  return 0;
}

int mjrl_peer_allreduce_f64(mjrl_ctx* ctx, mjrl_peer* p, const double* src, double* dst, int64_t count,
                            int64_t epoch, void* stream) {
  MJRL_REQUIRE(ctx && p && dst && p->connected && epoch >= 1, MJRL_E_ARG, "peer_allreduce: bad argument");
  MJRL_REQUIRE(count > 0 && count <= p->max_count, MJRL_E_SHAPE, "peer_allreduce: count %lld exceeds the block (%lld)",
               (long long)count, p->max_count);
  int rc = peer_kernel_attrs();
  if (rc) return rc;
  peer::allreduce_kernel<<<1, peer::THREADS, peer::BUF_ELEMS * sizeof(double), (cudaStream_t)stream>>>(
      p->view, src, dst, count, (unsigned long long)epoch);
  return mjrl_check_launch("peer allreduce");
}

int mjrl_cg_update_peer(mjrl_ctx* ctx, mjrl_peer* p, const mjrl_cg_state* st, int64_t d, double inv_count,
                        double damping, double residual_tol, int64_t epoch, void* stream) {
  MJRL_REQUIRE(ctx && p && st && p->connected && epoch >= 1, MJRL_E_ARG, "cg_update_peer: bad argument");
  MJRL_REQUIRE(d > 0 && d <= p->max_count, MJRL_E_SHAPE, "cg_update_peer: d exceeds the block");
  int rc = peer_kernel_attrs();
  if (rc) return rc;
  const size_t smem = peer::BUF_ELEMS * sizeof(double);
  if (d <= (int64_t)peer::CG_E * peer::THREADS)
    peer::cg_update_kernel<true><<<1, peer::THREADS, smem, (cudaStream_t)stream>>>(
        p->view, *st, d, inv_count, damping, residual_tol, (unsigned long long)epoch);
  else
    peer::cg_update_kernel<false><<<1, peer::THREADS, smem, (cudaStream_t)stream>>>(
        p->view, *st, d, inv_count, damping, residual_tol, (unsigned long long)epoch);
  return mjrl_check_launch("cg_update_peer");
}

// cg_solve.py:3-22 as ONE host call: cg_init, then `iters` x (product -> partial reduce -> CG update [fused with
// the peer all-reduce]) enqueued back to back from C.  A Python caller pays one ctypes transition instead of
// 2 * iters + 1: at small batches (and under strong scaling) the update is bound by how fast the host can enqueue.
int mjrl_cg_solve(mjrl_ctx* ctx, const mjrl_policy_desc* desc, const float* params, const float* obs, int64_t N,
                  const mjrl_cg_state* st, const double* b, double b_scale, int32_t iters, double inv_count,
                  double damping, double residual_tol, float* cache, int32_t mode_first, int32_t mode_rest,
                  int32_t jac_fill_after_first, mjrl_peer* peer, int64_t epoch0, void* stream) {
  MJRL_REQUIRE(ctx && desc && params && obs && st && b && N > 0 && iters >= 0, MJRL_E_ARG, "cg_solve: bad argument");
  MJRL_REQUIRE(peer == nullptr || (peer->connected && epoch0 >= 0), MJRL_E_ARG, "cg_solve: peer not connected");
  const int64_t d = mjrl_policy_param_count(desc);
  MJRL_REQUIRE(d > 0, MJRL_E_SHAPE, "cg_solve: bad policy description");
  int rc = mjrl_cg_init(ctx, st, d, b, b_scale, stream);
  if (rc) return rc;
  for (int32_t k = 0; k < iters; ++k) {
    const int64_t epoch = epoch0 + k + 1;
    double* out = peer ? mjrl_peer_slot(peer, epoch) : st->z;
    rc = mjrl_policy_fvp_cached(ctx, desc, params, obs, N, st->p32, out, st->flags, cache,
                                k == 0 ? mode_first : mode_rest, stream);
    if (rc) return rc;
    if (k == 0 && jac_fill_after_first) {
      rc = jac_fill_after_first == 2 ? mjrl_policy_fvp_tc_fill(ctx, desc, N, cache, stream)
                                     : mjrl_policy_fvp_jac_fill(ctx, desc, params, N, cache, stream);
      if (rc) return rc;
    }
    rc = peer ? mjrl_cg_update_peer(ctx, peer, st, d, inv_count, damping, residual_tol, epoch, stream)
              : mjrl_cg_update(ctx, st, d, inv_count, damping, residual_tol, stream);
    if (rc) return rc;
  }
  return 0;
}

int mjrl_peer_status(mjrl_peer* p, int32_t* err_out) {
  MJRL_REQUIRE(p && err_out, MJRL_E_ARG, "peer_status: bad argument");
  cudaSetDevice(p->device);
  int v = 0;
  cudaError_t e = cudaMemcpy(&v, p->err, sizeof(int), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { mjrl_set_error("peer_status: %s", cudaGetErrorString(e)); return (int)e; }
  *err_out = v;
  return 0;
}

int mjrl_peer_stamps(mjrl_peer* p, uint64_t* out, int64_t first_epoch, int64_t count) {
  MJRL_REQUIRE(p && out && first_epoch >= 0 && count >= 0 && count <= peer::STAMP_EPOCHS, MJRL_E_ARG, "peer_stamps: bad argument");
  MJRL_REQUIRE(p->stamps, MJRL_E_STATE, "peer_stamps: the block was created without MJRL_B200_PEER_STAMPS=1");
  cudaSetDevice(p->device);
  for (int64_t k = 0; k < count; ++k) {
    const int64_t e = (first_epoch + k) % peer::STAMP_EPOCHS;
    cudaError_t er = cudaMemcpy(out + 4 * k, p->stamps + 4 * e, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost);
    if (er != cudaSuccess) { mjrl_set_error("peer_stamps: %s", cudaGetErrorString(er)); return (int)er; }
  }
  return 0;
}

int mjrl_peer_destroy(mjrl_peer* p) {
  if (!p) return 0;
  cudaSetDevice(p->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < p->world; ++r)
    if (p->opened[r]) cudaIpcCloseMemHandle(p->bases[r]);
  if (p->base) cudaFree(p->base);
  if (p->err) cudaFree(p->err);
  if (p->stamps) cudaFree(p->stamps);
  delete p;
  return 0;
}

}  // extern "C"
