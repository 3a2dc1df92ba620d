// The checkpointed REVERSE SWEEP (north_star; pyhgf differentiates its scan with jax.grad, pyhgf/model/network.py:612-624).
// One thread carries one instance, as in hgf_scan.cuh, but through the time loop TWICE inside the same kernel:
//   1. forward over t = 0 .. T-1 with the generated straight-line step (csrc/generated/reverse_<topology>.h, written by
//      pyhgf_b200/revgen.py), summing the response surprise and storing the CARRIED beliefs of every step -- the beliefs
//      a step reads before writing them: 2 floats per continuous node -- into the checkpoint array state[T][kCarried][N]
//      (coalesced over instances);
//   2. backward over t = T-1 .. 0: reload step t's carried beliefs, run the step forward once more while logging the
//      previous value of every attribute slot each operation overwrites, seed the adjoint of the step's surprise, then
//      undo the operations one by one: restore the slots, recompute that ONE operation's intermediates in registers and
//      run its adjoint statements.  The adjoint attribute vector `sb` carries lambda_t to step t-1 through the carried
//      slots and accumulates d logp / d attribute in every slot no operation writes (the parameters).
// Cost: 3.6-3.9 plain filters measured (profiles/r2_grad_scaling.log), whatever the number of parameters; the attribute
// vector, its adjoint and the undo log of one step live in registers (every index is a literal after inlining).
#pragma once
#ifdef __CUDACC__
#include "hgf_dispatch.h"  // scratch_pool
#endif
#include "hgf_scan.cuh"

namespace hgf {

template <class F, class Rev, bool UNIT_DT>
HGF_D F reverse_instance(const Program& prog, const CallLists& lists, const ScanArgs& args, long long n, bool active, F* state) {
  const long long N = args.ld;
  F s[Rev::kSlots];
  const F* base = static_cast<const F*>(args.attributes);
#pragma unroll
  for (int k = 0; k < Rev::kSlots; ++k) s[k] = HGF_LDG(base + k);
  const F* params = static_cast<const F*>(args.params);
  for (int q = 0; q < lists.n_params; ++q) {
    const F v = HGF_LDG(params + static_cast<long long>(lists.param_row[q]) * N + n);
    const int slot = lists.param_slot[q];
#pragma unroll
    for (int k = 0; k < Rev::kSlots; ++k)
      if (k == slot) s[k] = v;
  }
  const int T = args.n_steps;
  const long long in_series = args.input_group_size > 0 ? n / args.input_group_size : 0;
  const F* u = static_cast<const F*>(args.input_data) + in_series * T * Rev::kInputs;
  const uint8_t* obs = args.observed ? args.observed + in_series * T * Rev::kInputs : nullptr;
  const F* resp = nullptr;
  if (args.response_inputs) {
    const long long rs = args.response_group_size > 0 ? n / args.response_group_size : 0;
    resp = static_cast<const F*>(args.response_inputs) + rs * T;
  }
  const F* time_steps = static_cast<const F*>(args.time_steps);
  const F cap = static_cast<F>(prog.max_posterior_precision);
  const int fn = args.response_fn;
  const long long plane = static_cast<long long>(Rev::kCarried) * N;

  auto observe = [&](int t) {
#pragma unroll
    for (int k = 0; k < Rev::kInputs; ++k) {
      const int node = Rev::input_node(k);
      s[node * NF + F_MEAN] = HGF_LDG(u + static_cast<long long>(t) * Rev::kInputs + k);
      s[node * NF + F_OBSERVED] = obs ? static_cast<F>(HGF_LDG(obs + static_cast<long long>(t) * Rev::kInputs + k)) : F(1);
    }
  };
  auto step_size = [&](int t, F& dt, F& log_dt) {
    if constexpr (UNIT_DT) { dt = F(1); log_dt = F(0); }
    else { dt = time_steps ? HGF_LDG(time_steps + t) : F(1); log_dt = b_log(dt); }
  };

  // ---- 1. forward, checkpointing the carried beliefs
  F total = F(0);
  bool node1_mean_nan = false;  // total_gaussian_surprise's any-NaN rule (response.py:86-89), as in hgf_scan.cuh
  for (int t = 0; t < T; ++t) {
#pragma unroll
    for (int c = 0; c < Rev::kCarried; ++c) state[t * plane + c * N + n] = s[Rev::carried_slot(c)];
    F dt, log_dt;
    step_size(t, dt, log_dt);
    Rev::predict(s, dt, log_dt, cap);
    observe(t);
    Rev::update(s, dt, log_dt, cap);
    total = total + Rev::response(fn, s, resp ? HGF_LDG(resp + t) : F(0));
    if (fn == RESP_TOTAL_GAUSSIAN && Rev::kNodes > 1 && b_isnan(s[1 * NF + F_MEAN])) node1_mean_nan = true;
  }

  // ---- 2. backward
  F sb[Rev::kSlots];
#pragma unroll
  for (int k = 0; k < Rev::kSlots; ++k) sb[k] = F(0);
  F log[Rev::kLog > 0 ? Rev::kLog : 1];
  for (int t = T - 1; t >= 0; --t) {
#pragma unroll
    for (int c = 0; c < Rev::kCarried; ++c) s[Rev::carried_slot(c)] = state[t * plane + c * N + n];
    F dt, log_dt;
    step_size(t, dt, log_dt);
    Rev::predict_log(s, dt, log_dt, cap, log);
    observe(t);
    Rev::update_log(s, dt, log_dt, cap, log);
    Rev::response_reverse(fn, s, resp ? HGF_LDG(resp + t) : F(0), F(1), sb);
    Rev::update_reverse(s, sb, log, dt, log_dt, cap);
#pragma unroll
    for (int k = 0; k < Rev::kInputs; ++k) {  // the observation overwrote these: nothing flows to before it
      sb[Rev::input_node(k) * NF + F_MEAN] = F(0);
      sb[Rev::input_node(k) * NF + F_OBSERVED] = F(0);
    }
    Rev::predict_reverse(s, sb, log, dt, log_dt, cap);
  }

  const F scale = static_cast<F>(args.out_scale);
  if (node1_mean_nan) total = b_inf<F>();  // the selected branch of the reference's jnp.where is a constant: zero gradient
  if (active) {
    if (args.surprise_sum) static_cast<F*>(args.surprise_sum)[n] = scale * total;
    if (args.gradient) {
      F* grad = static_cast<F*>(args.gradient);
      for (int q = 0; q < lists.n_params; ++q) {
        const int slot = lists.param_slot[q];
        F g = F(0);
#pragma unroll
        for (int k = 0; k < Rev::kSlots; ++k)
          if (k == slot) g = sb[k];
        grad[static_cast<long long>(lists.param_row[q]) * N + n] = node1_mean_nan ? F(0) : scale * g;
      }
    }
  }
  return total;
}

#ifdef __CUDACC__
// float32: bounded to 7 resident blocks per SM (128 registers; ptxas spills 30-400 bytes of the undo log) so that 65 536
// instances are one wave, like the forward-mode kernels (hgf_scan.cuh::scan_min_blocks): 3.8 -> 3.3 ms on config 4's batch
template <class F, class Rev, bool UNIT_DT>
__global__ void __launch_bounds__(SCAN_THREADS, (sizeof(F) == 4 && Rev::kNodes <= 4) ? 7 : 1)
reverse_kernel(const __grid_constant__ Program prog, const __grid_constant__ CallLists lists, const __grid_constant__ ScanArgs args, F* state) {
  long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool active = n < args.n_instances;
  if (!active) n = args.n_instances - 1;  // (its checkpoint column is rewritten with the same values)
  const F total = reverse_instance<F, Rev, UNIT_DT>(prog, lists, args, n, active, state);
  if (args.reduced) {
    const int g = args.group ? __ldg(args.group + n) : 0;
    reduce_into_groups<F>(static_cast<F*>(args.reduced), g, static_cast<F>(args.out_scale) * total, active);
  }
}

// Launch: the checkpoint array is stream-ordered scratch from the library's own pool (hgf_select.cu::scratch_pool;
// cudaMallocFromPoolAsync / cudaFreeAsync: no synchronisation, and the pool keeps the memory between launches).
template <class F, class Rev>
cudaError_t launch_reverse(const Program& prog, const CallLists& lists, const ScanArgs& args, cudaStream_t stream) {
  const long long blocks = (args.n_instances + SCAN_THREADS - 1) / SCAN_THREADS;
  if (blocks == 0) return cudaSuccess;
  F* state = nullptr;
  const size_t bytes = static_cast<size_t>(args.n_steps > 0 ? args.n_steps : 1) * Rev::kCarried * static_cast<size_t>(args.ld) * sizeof(F);
  cudaMemPool_t pool = nullptr;
  cudaError_t e = scratch_pool(&pool);
  if (e != cudaSuccess) return e;
  e = cudaMallocFromPoolAsync(reinterpret_cast<void**>(&state), bytes, pool, stream);
  if (e != cudaSuccess) return e;
  if (args.time_steps) reverse_kernel<F, Rev, false><<<static_cast<unsigned>(blocks), SCAN_THREADS, 0, stream>>>(prog, lists, args, state);
  else reverse_kernel<F, Rev, true><<<static_cast<unsigned>(blocks), SCAN_THREADS, 0, stream>>>(prog, lists, args, state);
  e = cudaGetLastError();
  const cudaError_t e2 = cudaFreeAsync(state, stream);
  return e != cudaSuccess ? e : e2;
}
#endif

}  // namespace hgf
