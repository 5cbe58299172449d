#pragma once
#include <climits>

#include "common.cuh"

namespace zb {

struct SamplerArgs {
  const float* logits;  // [B, Q, V] CFG-mixed, pad slot already -inf
  const float* noise;   // optional injected Exp(1) noise [n_draws][B, Q, V] (tests); nullptr -> Philox
  int32_t* delayed;     // [B, Q, t_delayed] delayed code matrix (-1 = not generated yet)
  int32_t *tok1, *tok2; // [B, Q] tokens of pass 1 / pass 2
  int* state;           // StepStateIdx
  int *remaining, *grace;                          // [B]
  uint8_t *stopping, *eos_mode, *mask2, *eos_old;  // [B]
  volatile int* host_flags;                        // pinned host mirror {done, steps} or nullptr
  int batch, n_q, vocab, t_delayed;
  int loop_mode;     // 0: first frame after prefill (no bias / penalty / state machine); 1: decode loop
  int first_offset;  // column written when loop_mode == 0
  int row_base;      // global index of utterance 0 (RNG stream id = (row_base + b) * Q + q)
  unsigned long long seed;
  float temperature, top_p, min_p, linear, conf, quad, rep_penalty;
  int top_k, rep_window;
};

cudaError_t launch_sampler(const SamplerArgs& a, int pass, cudaStream_t st);

}  // namespace zb
