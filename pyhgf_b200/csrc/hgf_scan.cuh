// The time loop: one thread carries one network instance through all T
// observations (Network.input_data's lax.scan, pyhgf/model/network.py:612-624,
// over beliefs_propagation, pyhgf/utils/beliefs_propagation.py:24-159), with
// the response function (pyhgf/response.py) and -- for Dual scalars -- its
// parameter gradient fused in.  Instances are independent, so the grid is a
// flat partition of the batch; every global access is coalesced over instances
// (parameters [P][N], trajectories [R][T][N], outputs [.][N]); the observation
// series is uniform across a warp (broadcast loads).
//
// Template parameters:
//   T       Dual<B, PF>: the filter's scalar (PF tangent directions; 0 = plain filter)
//   PR      extra tangent directions that exist only inside the response function
//           (response-function parameters cannot influence the beliefs)
//   UNIT_DT the call has no time_steps vector: dt = 1, log dt = 0 fold away
//   REC     the call records trajectories (the logp-only kernels carry no recording code at all)
#pragma once
#include "hgf_ops.cuh"

namespace hgf {

constexpr int SCAN_THREADS = 64;

template <class T, int PR> struct response_scalar { using type = Dual<base_t<T>, scalar_traits<T>::n_tangents + PR>; };
template <class T, int PR> using response_scalar_t = typename response_scalar<T, PR>::type;

// Philox4x32-10 (Salmon et al. 2011): the counter-based stream of the generative mode.  One call gives the four
// 32-bit words of counter (instance, step) under key `seed`.
HGF_HD void philox4x32(unsigned long long seed, unsigned long long instance, unsigned step, unsigned out[4]) {
  unsigned c0 = static_cast<unsigned>(instance), c1 = static_cast<unsigned>(instance >> 32), c2 = step, c3 = 0x48474600u;
  unsigned k0 = static_cast<unsigned>(seed), k1 = static_cast<unsigned>(seed >> 32);
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    const unsigned long long p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
    const unsigned n0 = static_cast<unsigned>(p1 >> 32) ^ c1 ^ k0, n1 = static_cast<unsigned>(p1);
    const unsigned n2 = static_cast<unsigned>(p0 >> 32) ^ c3 ^ k1, n3 = static_cast<unsigned>(p0);
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// One uniform in (0, 1) and one standard normal (Box-Muller) per instance-step, in double so that float32 and
// float64 launches see the same draws.
struct UniformNormal { double u, z; };
#ifdef __CUDACC__
static __host__ __device__ __noinline__  // a cold call: keeps its double-precision log / cos out of the time loop's registers
#else
static inline
#endif
UniformNormal draw_uniform_normal(unsigned long long seed, unsigned long long instance, unsigned step) {
  unsigned r[4];
  philox4x32(seed, instance, step, r);
  UniformNormal d;
  d.u = (static_cast<double>(r[0]) + 0.5) * (1.0 / 4294967296.0);
  const double u1 = (static_cast<double>(r[1]) + 0.5) * (1.0 / 4294967296.0);
  const double u2 = (static_cast<double>(r[2]) + 0.5) * (1.0 / 4294967296.0);
  d.z = ::sqrt(-2.0 * ::log(u1)) * ::cos(6.283185307179586 * u2);
  return d;
}

#ifdef __CUDA_ARCH__
// One TMA tensor store per block per step: the staged [rows][SCAN_THREADS] tile -> trajectories[0:rows][t][n0:n0+64]
// (cp.async.bulk.tensor, bulk-group completion; instances past N are clipped by the tensor map's bounds).
// Called by every thread of the block, in uniform control flow.
template <class B> __device__ __forceinline__ void tma_store_step(const ScanArgs& args, B* stage, int t) {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // this thread's staged values -> visible to the async proxy
  // the store of step t-1 must have finished reading its buffer before step t+1 overwrites it
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const B* src = stage + static_cast<size_t>(t & 1) * args.traj_rows * SCAN_THREADS;
    const unsigned smem = static_cast<unsigned>(__cvta_generic_to_shared(src));
    const int c0 = static_cast<int>(blockIdx.x) * SCAN_THREADS;
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(reinterpret_cast<unsigned long long>(args.traj_map)), "r"(c0), "r"(t), "r"(0), "r"(smem)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
}
#endif

// One instance, all time steps.  Returns the summed surprise (with its tangents);
// every per-instance output is stored here when `active`.
template <class T, class Topo, int NSLOTS, int PR, bool UNIT_DT, bool REC = true>
HGF_D response_scalar_t<T, PR> scan_instance(const Program& prog, const CallLists& lists, const ScanArgs& args, long long n,
                                             bool active, base_t<T>* stage = nullptr) {
  using B = base_t<T>;
  using TR = response_scalar_t<T, PR>;
  constexpr int PF = scalar_traits<T>::n_tangents;
  const Topo a(prog);
  const long long N = args.ld;  // row stride of every [.][N] array (== n_instances unless the launch covers a slice of a batch)

  // ---- initial attributes: shared vector + this instance's parameters ----
  T s[NSLOTS];
  const B* base = static_cast<const B*>(args.attributes);
  const int n_slots = a.n_slots();
  if constexpr (Topo::is_static) {
#pragma unroll
    for (int k = 0; k < NSLOTS; ++k)
      if (k < n_slots) s[k] = T(HGF_LDG(base + k));
  } else {
    for (int k = 0; k < n_slots; ++k) s[k] = T(HGF_LDG(base + k));
  }
  const B* params = static_cast<const B*>(args.params);
  for (int q = 0; q < lists.n_filter; ++q) {
    T v = T(HGF_LDG(params + static_cast<long long>(lists.param_row[q]) * N + n));
    if constexpr (PF > 0) {
#pragma unroll
      for (int i = 0; i < PF; ++i)
        if (i == q - args.tangent_offset) v.d.set(i, B(1));
    }
    const int slot = lists.param_slot[q];
    if constexpr (Topo::is_static) {
#pragma unroll
      for (int k = 0; k < NSLOTS; ++k)
        if (k == slot) s[k] = v;
    } else {
      s[slot] = v;
    }
  }
  // response-function parameters (the trailing entries of the lists override the shared values)
  TR rp[N_RESP_PARAMS];
#pragma unroll
  for (int j = 0; j < N_RESP_PARAMS; ++j) rp[j] = widen<TR>(s[a.slot_resp() + j]);
  for (int q = lists.n_filter; q < lists.n_params; ++q) {
    TR v = TR(HGF_LDG(params + static_cast<long long>(lists.param_row[q]) * N + n));
    if constexpr (PR > 0) {
      if (q == args.resp_index) v.d.set(PF, B(1));
    }
    const int j = lists.param_slot[q] - a.slot_resp();
#pragma unroll
    for (int jj = 0; jj < N_RESP_PARAMS; ++jj)
      if (jj == j) rp[jj] = v;
  }

  const int T_steps = args.n_steps;
  const int n_fed = a.n_fed();
  const long long in_series = args.input_group_size > 0 ? n / args.input_group_size : 0;
  const B* u = static_cast<const B*>(args.input_data) + in_series * T_steps * n_fed;
  const uint8_t* obs = args.observed ? args.observed + in_series * T_steps * n_fed : nullptr;
  const B* resp = nullptr;
  if (args.response_inputs) {
    const long long rs = args.response_group_size > 0 ? n / args.response_group_size : 0;
    resp = static_cast<const B*>(args.response_inputs) + rs * T_steps;
  }
  const B* time_steps = static_cast<const B*>(args.time_steps);
  B* traj = static_cast<B*>(args.trajectories);
  B* pointwise = static_cast<B*>(args.surprise_pointwise);
  const int response_fn = args.response_fn;

  TR total = TR(B(0));
  bool node1_mean_nan = false;
  // the static leaf set in slot order: f(leaf number, value)
  auto for_each_leaf = [&](auto&& f) {
    int leaf = 0;
#pragma unroll
    for (int node = 0; node < a.n_nodes(); ++node) {
#pragma unroll
      for (int fld = 0; fld < 4; ++fld, ++leaf) f(leaf, pv(HGF_S(node, fld)));
      if (a.node_type(node) == NODE_VOLATILE) {
#pragma unroll
        for (int fld = F_MEAN_VOL; fld < F_MEAN_VOL + 4; ++fld, ++leaf) f(leaf, pv(HGF_S(node, fld)));
      }
    }
  };
  // EVERY leaf recorded (config 5): row = leaf number, so the step's stores need no mask test, no branch and no computed offset
  bool rec_all = false;
  (void)for_each_leaf;
  if constexpr (REC && Topo::is_static) {
    int n_leaves = 0;
#pragma unroll
    for (int node = 0; node < a.n_nodes(); ++node) n_leaves += a.node_type(node) == NODE_VOLATILE ? 8 : 4;
    rec_all = n_leaves <= 128;  // (the mask holds 128 leaves: a larger topology with its first 128 recorded is NOT "all")
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int bits = n_leaves - 32 * w;
      const uint32_t full = bits >= 32 ? 0xffffffffu : (bits <= 0 ? 0u : (1u << bits) - 1u);
      rec_all = rec_all && args.rec_mask[w] == full;
    }
  }

  for (int t = 0; t < T_steps; ++t) {
    Step<T> st;
    if constexpr (UNIT_DT) {
      st.dt = B(1);
      st.log_dt = B(0);
    } else {
      st.dt = time_steps ? HGF_LDG(time_steps + t) : B(1);
      st.log_dt = x_log(st.dt);
    }
    // prediction sequence (parents before children)
    [[maybe_unused]] int pred_lo = 0, pred_hi = a.n_pred(), upd_lo = 0, upd_hi = a.n_upd();
    if constexpr (Topo::is_static) {
      a.run_predictions(a, s, st);
    } else {
      // a switching plan (input_custom_sequence, network.py:746-758) picks this step's branch; lax.switch clamps
      if (args.branch_idx) {
        int b = HGF_LDG(args.branch_idx + t);
        b = b < 0 ? 0 : (b >= a.n_branches() ? a.n_branches() - 1 : b);
        pred_lo = a.br_pred(b), pred_hi = a.br_pred(b + 1), upd_lo = a.br_upd(b), upd_hi = a.br_upd(b + 1);
      }
      for (int i = pred_lo; i < pred_hi; ++i) apply_op(a, s, st, a.pred_op(i), a.pred_node(i));
    }
    // set_observation, pyhgf/updates/observation.py:10-38
    if (REC && args.generative) {  // rollouts always record: only the recording kernels carry this branch
      // observations = "generative" (beliefs_propagation.py:109-125): every input node observes a draw from its own
      // prediction (sample_node_distribution.py:34-40); one draw per step shared by the input nodes, as the reference
      // samples all of them with the same key
      const UniformNormal draw = draw_uniform_normal(args.seed, static_cast<unsigned long long>(n), static_cast<unsigned>(t));
      const double du = draw.u, dz = draw.z;
#pragma unroll (Topo::kUnroll)
      for (int k = 0; k < a.n_inputs(); ++k) {
        const int node = a.input_idx(k);
        const B mu = pv(HGF_S(node, F_EXPECTED_MEAN));
        B value;
        if (a.node_type(node) == NODE_BINARY) value = (static_cast<B>(du) < mu) ? B(1) : B(0);
        else value = static_cast<B>(dz) * (B(1) / static_cast<B>(::sqrt(static_cast<double>(pv(HGF_S(node, F_EXPECTED_PRECISION)))))) + mu;
        HGF_S(node, F_MEAN) = T(value);
        HGF_S(node, F_OBSERVED) = T(B(1));
      }
    } else {
#pragma unroll (Topo::kUnroll)
      for (int k = 0; k < n_fed; ++k) {
        const int node = a.input_idx(k);
        HGF_S(node, F_MEAN) = T(HGF_LDG(u + static_cast<long long>(t) * n_fed + k));
        HGF_S(node, F_OBSERVED) = T(obs ? static_cast<B>(HGF_LDG(obs + static_cast<long long>(t) * n_fed + k)) : B(1));
      }
    }
    // prediction errors and posterior updates (children before parents)
    if constexpr (Topo::is_static) {
      a.run_updates(a, s, st);
    } else {
      for (int i = upd_lo; i < upd_hi; ++i) apply_op(a, s, st, a.upd_op(i), a.upd_node(i));
    }

    if (response_fn != RESP_NONE) {
      const TR r = response_step<TR>(a, s, rp, response_fn, resp ? HGF_LDG(resp + t) : B(0));
      total = total + r;
      if (pointwise && active) pointwise[static_cast<long long>(t) * N + n] = pv(r);
      if (response_fn == RESP_TOTAL_GAUSSIAN && a.n_nodes() > 1 && x_isnan(HGF_S(1, F_MEAN))) node1_mean_nan = true;
    }
    if (REC && traj && (active || stage)) {
      // `stage` (TMA write-back): this thread's column of the block's [rows][SCAN_THREADS] staging tile for step t
      // (double-buffered); otherwise a direct store, coalesced over the warp's instances
      B* col = nullptr;
#ifdef __CUDA_ARCH__
      if (stage) col = stage + static_cast<size_t>(t & 1) * args.traj_rows * SCAN_THREADS + threadIdx.x;
#endif
      B* direct = traj + static_cast<long long>(t) * N + n;
      const long long leaf_stride = static_cast<long long>(T_steps) * N;
      if constexpr (Topo::is_static) {
        // static leaf set (see ScanArgs::rec_mask): warp-uniform predicates
        if (rec_all) {
          if (col) for_each_leaf([&](int leaf, B v) { col[leaf * SCAN_THREADS] = v; });
          else for_each_leaf([&](int leaf, B v) { direct[leaf * leaf_stride] = v; });
        } else {
          int row = 0;
          for_each_leaf([&](int leaf, B v) {
            if (args.rec_mask[leaf >> 5] >> (leaf & 31) & 1u) {
              if (col) col[row * SCAN_THREADS] = v; else direct[row * leaf_stride] = v;
              ++row;
            }
          });
        }
      } else {
        for (int r = 0; r < lists.n_rec; ++r) {
          if (col) col[r * SCAN_THREADS] = pv(s[lists.rec_slot[r]]); else direct[r * leaf_stride] = pv(s[lists.rec_slot[r]]);
        }
      }
#ifdef __CUDA_ARCH__
      if (stage) tma_store_step(args, stage, t);
#endif
    }
  }
#ifdef __CUDA_ARCH__
  if (REC && traj && stage && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // smem outlives the last store
#endif

  // response.py:86-89: total_gaussian_surprise is +inf at EVERY trial if node 1's mean ever was NaN -- the sum, and
  // the per-trial values already written (a rare, cold rewrite of this instance's column)
  if (response_fn == RESP_TOTAL_GAUSSIAN && node1_mean_nan) {
    total = TR(b_inf<B>());
    if (pointwise && active)
      for (int t = 0; t < T_steps; ++t) pointwise[static_cast<long long>(t) * N + n] = b_inf<B>();
  }

  const B scale = static_cast<B>(args.out_scale);
  if (active) {
    if (args.surprise_sum) static_cast<B*>(args.surprise_sum)[n] = scale * pv(total);
    if (args.gradient) {
      B* grad = static_cast<B*>(args.gradient);
      if constexpr (PF > 0) {
#pragma unroll
        for (int i = 0; i < PF; ++i)
          if (args.tangent_offset + i < lists.n_filter)
            grad[static_cast<long long>(lists.param_row[args.tangent_offset + i]) * N + n] = scale * total.d.get(i);
      }
      if constexpr (PR > 0) {
        if (args.resp_index >= 0) grad[static_cast<long long>(lists.param_row[args.resp_index]) * N + n] = scale * total.d.get(PF);
      }
    }
    // the final attribute vector is an interpreter-only output (hgf_static_run.h routes such calls there):
    // keeping all n_nodes x 19 fields alive to the end of the loop would cost the register-resident
    // kernels their registers
    if constexpr (!Topo::is_static) {
      if (args.last_attributes) {
        B* last = static_cast<B*>(args.last_attributes);
        for (int k = 0; k < a.n_nodes() * NF; ++k) last[static_cast<long long>(k) * N + n] = pv(s[k]);
      }
    }
  }
  return total;
}

#ifdef __CUDACC__
// warp-aggregated atomic accumulation of per-instance sums into per-group sums
template <class B> __device__ __forceinline__ void reduce_into_groups(B* reduced, int g, B value, bool active) {
  const unsigned full = 0xffffffffu;
  const int g0 = __shfl_sync(full, g, 0);
  const bool uniform = __all_sync(full, g == g0);
  if (uniform) {
    B v = active ? value : B(0);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(full, v, off);
    if ((threadIdx.x & 31) == 0) atomicAdd(reduced + g0, v);
  } else if (active) {
    atomicAdd(reduced + g, value);
  }
}

// ---------------------------------------------------------------------------
// Fused all-reduce of the per-group sums over NVLink peer memory (hgf_call_t::peer_workspace; layout and protocol:
// hgf_program.h).  Every block has added its groups' sums into this rank's `partial`.  The LAST block of the grid
// (ticket counter) then
//   (0) flow control: makes sure no peer is more than FUSED_RING / 2 - 1 launches behind (it normally is not, and
//       nothing spins) -- that is what guarantees that the slot written below was zeroed by its owner;
//   (1) zeroes this rank's slot of launch e + FUSED_RING / 2;
//   (2) adds each partial sum into slot e % FUSED_RING of EVERY rank -- one multimem.red through the NVSwitch
//       multicast mapping, or one system-scope red per peer;
//   (3) fences and stores e + 1 into its `done` word on every rank;
// and the kernel ends.  It does NOT wait for the other ranks' sums: a rank that runs ahead is not held back by the
// launch skew of the others.  The consumer of the sums waits (fused_wait_kernel, hgf_fused_reduce_wait).
// Called by every thread of every block, in uniform control flow.
template <class B> __device__ __forceinline__ void red_add_sys(B* p, B v) { atomicAdd_system(p, v); }
__device__ __forceinline__ void multimem_red(float* p, float v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_red(double* p, double v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Spin until done[p] >= want (wrap-safe); false after ~10 s: a peer never arrived.
__device__ __forceinline__ bool fused_wait_done(const unsigned* done_word, unsigned want) {
  const long long t0 = clock64();
  while (static_cast<int>(ld_acquire_sys(done_word) - want) < 0) {
    if (clock64() - t0 > 20000000000LL) return false;
  }
  return true;
}

template <class B> __device__ __forceinline__ void fused_deposit(const ScanArgs& args) {
  __shared__ int is_last;
  const int G = args.n_groups, W = args.world;
  unsigned char* base = static_cast<unsigned char*>(args.peer_ws[args.rank]);
  B* ws = reinterpret_cast<B*>(base);
  B* partial = ws + static_cast<size_t>(FUSED_RING) * G;
  unsigned* ticket = reinterpret_cast<unsigned*>(base + fused_ctl_offset(G, sizeof(B)));
  const unsigned* done = reinterpret_cast<const unsigned*>(base + fused_done_offset(G, sizeof(B)));
  __syncthreads();  // this block's atomics into `partial` are issued
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const unsigned e = args.epoch;
  // (0) every peer has finished launch e - FUSED_RING / 2 (and with it zeroed its slot e % FUSED_RING)
  bool ok = true;
  if (e >= FUSED_RING / 2 && static_cast<int>(threadIdx.x) < W) ok = fused_wait_done(done + threadIdx.x, e - FUSED_RING / 2 + 1u);
  ok = __syncthreads_and(ok);
  // (1)
  B* ahead = ws + static_cast<size_t>((e + FUSED_RING / 2) % FUSED_RING) * G;
  for (int g = threadIdx.x; g < G; g += blockDim.x) ahead[g] = B(0);
  // (2)
  const size_t cur = static_cast<size_t>(e % FUSED_RING) * G;
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    const B v = ok ? __ldcg(partial + g) : b_nan<B>();  // a peer that never arrived: poison instead of hanging the GPU
    partial[g] = B(0);  // ready for the next launch
    if (args.mc_ws) {
      multimem_red(static_cast<B*>(args.mc_ws) + cur + g, v);
    } else {
      for (int p = 0; p < W; ++p) red_add_sys(static_cast<B*>(args.peer_ws[p]) + cur + g, v);
    }
  }
  // (3)
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) *ticket = 0u;  // for the next launch
  if (static_cast<int>(threadIdx.x) < W) {
    unsigned* peer_done = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(args.peer_ws[threadIdx.x]) + fused_done_offset(G, sizeof(B)));
    st_release_sys(peer_done + args.rank, e + 1u);
  }
}

// The consumer's side: wait until every rank has deposited launch `epoch` into this rank's workspace.  One block;
// thread p watches rank p.  A peer that never arrives poisons the slot with NaN after ~10 s.
template <class B> __global__ void fused_wait_kernel(void* workspace, int G, int W, unsigned epoch) {
  unsigned char* base = static_cast<unsigned char*>(workspace);
  const unsigned* done = reinterpret_cast<const unsigned*>(base + fused_done_offset(G, sizeof(B)));
  bool ok = true;
  if (static_cast<int>(threadIdx.x) < W) ok = fused_wait_done(done + threadIdx.x, epoch + 1u);
  ok = __syncthreads_and(ok);
  if (!ok) {
    B* slot = reinterpret_cast<B*>(base) + static_cast<size_t>(epoch % FUSED_RING) * G;
    for (int g = threadIdx.x; g < G; g += blockDim.x) slot[g] = b_nan<B>();
  }
}

// Occupancy: BASELINE config 4 gives each GPU 65 536 instances = 13.8 warps per SM; the float32 register-
// resident kernels of the small topologies are bounded to 7 resident 64-thread blocks (<= 144 registers) so that the
// whole batch is co-resident in one wave instead of 1.7 (a second, mostly empty wave costs ~40 %).  Topologies of up to
// 8 nodes are bounded to 8 blocks (128 registers, no spills: ptxas -v): BASELINE config 5's 16-node network runs as its two
// 8-node components on two streams (Network.batch), 2 x 512 blocks for 32 768 instances, and at the 151 registers the
// compiler takes unbounded only 6 blocks fit an SM -- 888 slots for 1 024 blocks, i.e. a second wave that made the two
// launches take as long as one after the other (profiles/r2_ncu_counters.json: 0.755 ms each alone, 1.45 ms together).
#ifndef HGF_WIDE_MIN_BLOCKS
#define HGF_WIDE_MIN_BLOCKS 1  // resident 64-thread blocks per SM asked of the wide (> 8 nodes) float32 specialisations
#endif
#ifndef HGF_MID_MIN_BLOCKS
#define HGF_MID_MIN_BLOCKS 8  // the same for 5 .. 8 nodes (tools/bench_c5_blocks.py measures 7 and 6 against it)
#endif
template <class T, class Topo> constexpr int scan_min_blocks() {
  if constexpr (!Topo::is_static) return 1; else
  return sizeof(base_t<T>) != 4 ? 1 : (Topo::kNodes <= 4 ? 7 : (Topo::kNodes <= 8 ? HGF_MID_MIN_BLOCKS : HGF_WIDE_MIN_BLOCKS));
}

template <class T, class Topo, int NSLOTS, int PR, bool UNIT_DT, bool REC>
#ifdef HGF_SCAN_MAXNREG  // experiments only (tools/bench_c5_blocks.py): an explicit register cap instead of a residency bound
__global__ void __maxnreg__(HGF_SCAN_MAXNREG)
#else
__global__ void __launch_bounds__(SCAN_THREADS, scan_min_blocks<T, Topo>())
#endif
scan_kernel(const __grid_constant__ Program prog, const __grid_constant__ CallLists lists, const __grid_constant__ ScanArgs args) {
  using B = base_t<T>;
  long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool active = n < args.n_instances;
  if (!active) n = args.n_instances - 1;  // keep the warp whole for the shuffle reduction; stores are predicated
  extern __shared__ __align__(128) unsigned char hgf_stage[];
  B* stage = (REC && args.traj_rows > 0) ? reinterpret_cast<B*>(hgf_stage) : nullptr;
  const auto total = scan_instance<T, Topo, NSLOTS, PR, UNIT_DT, REC>(prog, lists, args, n, active, stage);
  if (args.world > 1) {
    const int g = args.group ? __ldg(args.group + n) : 0;
    B* partial = static_cast<B*>(args.peer_ws[args.rank]) + static_cast<size_t>(FUSED_RING) * args.n_groups;
    reduce_into_groups<B>(partial, g, static_cast<B>(args.out_scale) * pv(total), active);
    fused_deposit<B>(args);
  } else if (args.reduced) {
    const int g = args.group ? __ldg(args.group + n) : 0;
    reduce_into_groups<B>(static_cast<B*>(args.reduced), g, static_cast<B>(args.out_scale) * pv(total), active);
  }
}
#endif

}  // namespace hgf
