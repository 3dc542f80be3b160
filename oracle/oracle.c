/*
 * Plain-C restatement of the arithmetic on the reference's data-parallel hot path.
 *
 * TEST INFRASTRUCTURE ONLY: linked by nothing in the product package; loaded by tests/ through
 * oracle/c_oracle.py (and built by __graft_entry__.build()). Same parity status as
 * oracle/ref_numpy.py: the reference ships no tests or golden vectors for this path ("parity
 * unpinned" by the reference itself), so this file is pinned against the outputs of the image's
 * torch 2.11.0 / transformers 5.5.0 CPU ops committed under tests/golden/.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (no FMA contraction: the roundings below are
 * part of the specification).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* (a1/a2) label-masked cross-entropy, mean over rows with label != ignore_index.
 * reference: train-accelerator.py:220-222, train-task.py:289-292, train-torchrun.py:140
 * library  : SP/transformers/models/t5/modeling_t5.py:1112-1117 (CrossEntropyLoss(ignore_index=-100))
 * double-precision math, rounded once. dlogits may be NULL. Returns the number of valid rows. */
int64_t o_ce(const float* logits, const int64_t* labels, int64_t n, int64_t v, int64_t ignore_index,
             double grad_out, float* loss_out, float* row_lse, double* dlogits) {
  int64_t n_valid = 0;
  double total = 0.0;
  for (int64_t i = 0; i < n; ++i)
    if (labels[i] != ignore_index) ++n_valid;
  for (int64_t i = 0; i < n; ++i) {
    const float* z = logits + i * v;
    double m = -INFINITY;
    for (int64_t j = 0; j < v; ++j)
      if ((double)z[j] > m) m = (double)z[j];
    double s = 0.0;
    for (int64_t j = 0; j < v; ++j) s += exp((double)z[j] - m);
    const double lse = m + log(s);
    if (row_lse) row_lse[i] = (float)lse;
    const int valid = labels[i] != ignore_index;
    if (valid) total += lse - (double)z[labels[i]];
    if (dlogits) {
      double* d = dlogits + i * v;
      if (!valid || n_valid == 0) {
        memset(d, 0, (size_t)v * sizeof(double)); /* torch writes exact zeros for ignored rows */
      } else {
        const double c = grad_out / (double)n_valid;
        for (int64_t j = 0; j < v; ++j) d[j] = exp((double)z[j] - lse) * c;
        d[labels[i]] = (exp((double)z[labels[i]] - lse) - 1.0) * c;
      }
    }
  }
  *loss_out = n_valid ? (float)(total / (double)n_valid) : NAN;
  return n_valid;
}

/* (a5) AdamW step, fp32, operation order and roundings of torch's single-tensor CPU path
 * reference: train-accelerator.py:187,224; train-task.py:260,295; HF Trainer for train-torchrun.py:140
 * library  : SP/torch/optim/adam.py:436-536 (mul_ decay; lerp_ = fma; mul_ + addcmul_;
 *            sqrt / bias_correction2_sqrt + eps; addcdiv_). grad_scale multiplies g first. */
void o_adamw_step(float* p, const float* g, float* m, float* v, int64_t n, double lr, double beta1,
                  double beta2, double eps, double weight_decay, int64_t step, double grad_scale) {
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  const float decay = (float)(1.0 - lr * weight_decay);
  const float w1 = (float)(1.0 - beta1);
  const float b2 = (float)beta2;
  const float omb2 = (float)(1.0 - beta2);
  const float neg_step = (float)(-(lr / bc1));
  const float bc2_sqrt = (float)sqrt(bc2);
  const float epsf = (float)eps;
  const float gs = (float)grad_scale;
  for (int64_t i = 0; i < n; ++i) {
    const float gi = grad_scale == 1.0 ? g[i] : g[i] * gs;
    float pi = p[i] * decay;
    const float mi = fmaf(w1, gi - m[i], m[i]);
    const float t = omb2 * gi;
    const float vi = v[i] * b2 + t * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + epsf;
    pi = pi + (neg_step * mi) / denom;
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
  }
}

/* (a3/a4) gradient mean all-reduce: sum over ranks of scale * b_r, accumulated in rank order in fp32
 * reference: DDP Reducer (train-accelerator.py:189,222; train-torchrun.py:127,140), train-task.py:65-69
 * library  : default_hooks.py:18-54 (div_ by world, then SUM). bufs: world pointers to n floats. */
void o_allreduce_mean(const float* const* bufs, int world, int64_t n, float scale, float* out) {
  for (int64_t i = 0; i < n; ++i) {
    float acc = bufs[0][i] * scale;
    for (int r = 1; r < world; ++r) acc = acc + bufs[r][i] * scale;
    out[i] = acc;
  }
}

/* (f1) clip coefficient of torch.nn.utils.clip_grad_norm_ over several tensors, L2:
 * per-tensor fp32 norms are not reproduced; the norm is accumulated in double. */
double o_grad_norm(const float* const* grads, const int64_t* numels, int n_tensors) {
  double s = 0.0;
  for (int t = 0; t < n_tensors; ++t)
    for (int64_t i = 0; i < numels[t]; ++i) s += (double)grads[t][i] * (double)grads[t][i];
  return sqrt(s);
}

/* (a6) seq2seq collate, int64, right padding
 * reference: DataCollatorForSeq2Seq at train-torchrun.py:113, train-accelerator.py:155-159, train-task.py:242
 * library  : SP/transformers/data/data_collator.py:529-615 + T5._shift_right (modeling_t5.py:595-613)
 * rows are given flattened with offsets (CSR style); out arrays are [b, ls] / [b, lt]. */
void o_collate(const int64_t* ids, const int64_t* mask, const int64_t* id_off, const int64_t* lab,
               const int64_t* lab_off, int64_t b, int64_t ls, int64_t lt, int64_t pad_id,
               int64_t label_pad, int64_t decoder_start, int64_t* input_ids, int64_t* attention,
               int64_t* labels, int64_t* decoder_input_ids) {
  for (int64_t i = 0; i < b; ++i) {
    const int64_t n = id_off[i + 1] - id_off[i];
    for (int64_t j = 0; j < ls; ++j) {
      input_ids[i * ls + j] = j < n ? ids[id_off[i] + j] : pad_id;
      attention[i * ls + j] = j < n ? mask[id_off[i] + j] : 0;
    }
    const int64_t nl = lab_off[i + 1] - lab_off[i];
    for (int64_t j = 0; j < lt; ++j) labels[i * lt + j] = j < nl ? lab[lab_off[i] + j] : label_pad;
    for (int64_t j = 0; j < lt; ++j) {
      int64_t t = j == 0 ? decoder_start : labels[i * lt + j - 1];
      decoder_input_ids[i * lt + j] = t == -100 ? pad_id : t;
    }
  }
}
