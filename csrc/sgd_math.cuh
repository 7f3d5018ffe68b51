// The one definition of the parameter update, shared by the single-GPU kernel (optim.cu) and the
// fused all-reduce epilogue (comm.cu).  torch.optim.SGD semantics (torch/optim/sgd.py:343-380) with
// dampening 0: g += wd*p; buf = mu*buf + g (zero-initialised buf == torch's first-step buf = g);
// g = nesterov ? g + mu*buf : buf; p -= lr*g.  `grad_scale` carries the 1/world_size of the average.
#pragma once

namespace pmd {

struct SgdHyper {
  float lr, momentum, weight_decay, grad_scale;
  int nesterov;
};

__device__ __forceinline__ void sgd_update(float& p, float g, float& m, const SgdHyper& h) {
  g = fmaf(g, h.grad_scale, h.weight_decay * p);
  m = fmaf(h.momentum, m, g);
  const float upd = h.nesterov ? fmaf(h.momentum, m, g) : m;
  p = fmaf(-h.lr, upd, p);
}

}  // namespace pmd
