/* oracle/restate.c — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Plain-C, single-threaded CPU restatement of the reference's learner hot path
 * (cemlyn007/ale-libtorch-ppo).  Every function cites the reference file:line it follows.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 * Compiled with -ffp-contract=off so each float op is rounded on its own, the way the
 * reference's one-ATen-op-per-arithmetic-step code is.
 *
 * Parity pinning: checked against (1) the six known-answer GAE cases and the two vision
 * cases of the reference's own tests (tests/test_oracle_golden.py), (2) fixtures generated
 * by the reference's own code (oracle/_ref, tests/golden/), see DESIGN.md.
 *
 * Layouts are the reference's: rollout storage env-major [E,T,...]; bools 1 byte;
 * actions int64; observations [.,4,84,84] u8 (plane 0 newest); parameters flat in
 * `parameters()` order c1w c1b c2w c2b c3w c3b fcw fcb piw pib vw vb, each OIHW / [out,in].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define RS_OK 0
#define RS_INVALID_ARGUMENT 1
#define RS_RUNTIME_ERROR 2

/* ---- GAE: src/ai/gae.cc:49-79 ---------------------------------------------------- */
int rs_gae(float *adv, const float *rewards, const float *values, const float *next_values,
           const uint8_t *terminals, const uint8_t *truncations, const uint8_t *starts,
           int64_t E, int64_t T, float gamma, float lambda) {
  for (int64_t i = 0; i < E * T; ++i) /* gae.cc:49-53 exclusivity */
    if ((int)(starts[i] != 0) + (int)(terminals[i] != 0) + (int)(truncations[i] != 0) > 1)
      return RS_INVALID_ARGUMENT;
  const float gl = gamma * lambda; /* float*float product, gae.cc:63 */
  for (int64_t e = 0; e < E; ++e) {
    float last = 0.0f, nv = next_values[e];
    for (int64_t t = T - 1; t >= 0; --t) {
      const int64_t i = e * T + t;
      const float r = rewards[i], v = values[i];
      const float gnv = gamma * nv;
      float running = r + gnv;
      running = running - v;
      float carry = gl * last;
      running = running + carry;              /* gae.cc:61-63 */
      const float terminal = r - v;           /* gae.cc:64 */
      float trunc = r + gnv;
      trunc = trunc - v;                      /* gae.cc:65-66 */
      float a = starts[i] ? 0.0f : running;   /* gae.cc:68-69 */
      a = terminals[i] ? terminal : a;        /* gae.cc:70-71 */
      a = truncations[i] ? trunc : a;         /* gae.cc:72-73 */
      adv[i] = a;
      last = a;                               /* gae.cc:77 */
      nv = v;                                 /* gae.cc:78 */
    }
  }
  return RS_OK;
}

/* ---- Buffer::get: src/ai/buffer.cc:58-77 ------------------------------------------ */
int rs_buffer_get(float *rewards_inout, const float *values, const float *next_values,
                  const uint8_t *terminals, const uint8_t *truncations, const uint8_t *starts,
                  int64_t E, int64_t T, float gamma, float lambda, int64_t write_index,
                  float *adv, float *returns, uint8_t *masks) {
  if (write_index != 0)
    return RS_RUNTIME_ERROR; /* buffer.cc:64-65 */
  for (int64_t i = 0; i < E * T; ++i) { /* buffer.cc:67 clamp_ in place */
    float r = rewards_inout[i];
    rewards_inout[i] = r < -1.0f ? -1.0f : (r > 1.0f ? 1.0f : r);
  }
  int rc = rs_gae(adv, rewards_inout, values, next_values, terminals, truncations, starts, E,
                  T, gamma, lambda);
  if (rc)
    return rc;
  for (int64_t i = 0; i < E * T; ++i) {
    returns[i] = adv[i] + values[i]; /* buffer.cc:70-71 */
    masks[i] = starts[i] ? 0 : 1;    /* buffer.cc:74 */
  }
  return RS_OK;
}

/* ---- preprocess: max_and_skip.cc:33-39, vision.cc:8-10,38-48 (area), :51,:71-84 (gray) */
static int bin_start(int o, int in, int out) { return (o * in) / out; }            /* floor */
static int bin_end(int o, int in, int out) { return ((o + 1) * in + out - 1) / out; } /* ceil */

/* frames_*: [B,210,160,3] u8 HWC -> gray_f32 [B,84,84] (unquantised), gray_u8 = floor(x+.5) */
int rs_preprocess(const uint8_t *frames_a, const uint8_t *frames_b, int64_t B, float *gray_f32,
                  uint8_t *gray_u8) {
  const float w[3] = {0.2125f, 0.7154f, 0.0721f}; /* vision.cc:51 */
  for (int64_t b = 0; b < B; ++b) {
    const uint8_t *fa = frames_a + b * 100800, *fb = frames_b + b * 100800;
    for (int oy = 0; oy < 84; ++oy) {
      const int y0 = bin_start(oy, 210, 84), y1 = bin_end(oy, 210, 84);
      for (int ox = 0; ox < 84; ++ox) {
        const int x0 = bin_start(ox, 160, 84), x1 = bin_end(ox, 160, 84);
        float mean[3];
        for (int c = 0; c < 3; ++c) {
          float sum = 0.0f;
          for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
              const int i = (y * 160 + x) * 3 + c;
              const uint8_t m = fa[i] > fb[i] ? fa[i] : fb[i]; /* max_and_skip.cc:36-38 */
              sum += (float)m;
            }
          mean[c] = sum / (float)(y1 - y0) / (float)(x1 - x0); /* adaptive avg pool */
        }
        float g = w[0] * mean[0];
        g = g + w[1] * mean[1];
        g = g + w[2] * mean[2]; /* vision.cc:80 matmul over the channel axis */
        const int64_t o = (b * 84 + oy) * 84 + ox;
        if (gray_f32)
          gray_f32[o] = g;
        if (gray_u8) {
          float q = floorf(g + 0.5f);
          gray_u8[o] = (uint8_t)(q < 0.0f ? 0.0f : (q > 255.0f ? 255.0f : q));
        }
      }
    }
  }
  return RS_OK;
}

/* ---- frame stack: src/ai/rollout.cc:184-196. stack [E,4,84,84], plane 0 newest ------ */
int rs_stack_update(uint8_t *stack, const uint8_t *frames, const uint8_t *starts, int64_t E) {
  for (int64_t e = 0; e < E; ++e) {
    uint8_t *s = stack + e * 4 * 7056;
    const uint8_t *f = frames + e * 7056;
    if (starts[e]) {
      for (int p = 0; p < 4; ++p)
        memcpy(s + p * 7056, f, 7056); /* rollout.cc:189-193 */
    } else {
      memmove(s + 7056, s, 3 * 7056);  /* rollout.cc:185-188 */
      memcpy(s, f, 7056);              /* rollout.cc:194-195 */
    }
  }
  return RS_OK;
}

/* ---- losses: src/ai/ppo/losses.cc:4-47 (+ analytic backward of the autograd graph) ---- */
void rs_log_softmax(const float *logits, int64_t n, int64_t A, float *out) { /* losses.cc:45-47 */
  for (int64_t i = 0; i < n; ++i) {
    const float *z = logits + i * A;
    float mx = z[0];
    for (int64_t j = 1; j < A; ++j)
      mx = z[j] > mx ? z[j] : mx;
    float s = 0.0f;
    for (int64_t j = 0; j < A; ++j)
      s += expf(z[j] - mx);
    const float lse = mx + logf(s);
    for (int64_t j = 0; j < A; ++j)
      out[i * A + j] = z[j] - lse;
  }
}

/* logits are RAW network outputs; old_log_probs are normalised (prepare_batch, train.cc:279).
 * Outputs per sample: clipped (surrogate), value_losses, entropies, total_losses, ratio;
 * scalar loss = sum(mask*total)/sum(mask) (losses.cc:19); dlogits/dvalues = d loss / d(raw
 * logits, value).  inv_count_override > 0 replaces 1/sum(mask) (multi-GPU global count). */
/* old_values != NULL: the opt-in clipped value loss (NOT in the reference, losses.cc:15 is the plain
 * form; north_star / SURVEY D3): vl = 0.5 max((v-R)^2, (v_old + clamp(v-v_old, -clip, clip) - R)^2),
 * backward with torch semantics (max -> larger branch, ties 1/2 + 1/2; clamp passes on the closed range) */
int rs_loss_ex(const float *logits, const float *old_log_probs, const int64_t *actions,
               const float *advantages, const float *values, const float *returns,
               const uint8_t *masks, int64_t n, int64_t A, float clip, float vcoef, float ecoef,
               float *loss, float *clipped, float *value_losses, float *entropies,
               float *total_losses, float *ratio, float *dlogits, float *dvalues,
               const float *old_values) {
  float *logp = (float *)malloc(sizeof(float) * (size_t)(n * A));
  rs_log_softmax(logits, n, A, logp);
  double count = 0.0;
  for (int64_t i = 0; i < n; ++i)
    count += masks[i] ? 1.0 : 0.0;
  const float inv = (float)(1.0 / count);
  double acc = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    const float *lp = logp + i * A;
    const int64_t a = actions[i];
    const float adv = advantages[i];
    const float rho = expf(lp[a] - old_log_probs[i * A + a]);          /* losses.cc:33 */
    float crho = rho < 1.0f - clip ? 1.0f - clip : (rho > 1.0f + clip ? 1.0f + clip : rho);
    const float un = rho * adv, cl = crho * adv;                       /* losses.cc:36-37 */
    const float surr = un < cl ? un : cl;                              /* losses.cc:38 */
    const float dv = values[i] - returns[i];
    float vl = 0.5f * (dv * dv);                                       /* losses.cc:15 */
    float dvl = dv;
    if (old_values) {
      const float d0 = values[i] - old_values[i];
      const float cd = d0 < -clip ? -clip : (d0 > clip ? clip : d0);
      const float dc = old_values[i] + cd - returns[i];
      const float lu = dv * dv, lc = dc * dc;
      const float gc = (d0 >= -clip && d0 <= clip) ? dc : 0.0f;
      vl = 0.5f * (lu > lc ? lu : lc);
      dvl = lu > lc ? dv : (lu < lc ? gc : 0.5f * dv + 0.5f * gc);
    }
    float ent = 0.0f;
    for (int64_t j = 0; j < A; ++j)
      ent += expf(lp[j]) * lp[j];
    ent = -ent;                                                        /* losses.cc:41-43 */
    const float total = -surr + vcoef * vl - ecoef * ent;              /* losses.cc:17-18 */
    if (clipped) clipped[i] = surr;
    if (value_losses) value_losses[i] = vl;
    if (entropies) entropies[i] = ent;
    if (total_losses) total_losses[i] = total;
    if (ratio) ratio[i] = rho;
    if (masks[i])
      acc += (double)total;
    /* backward: torch.min splits ties 1/2+1/2; clamp passes gradient on the closed range */
    const int inside = rho >= 1.0f - clip && rho <= 1.0f + clip;
    const float g_rho = (inside || un < cl) ? adv : 0.0f;
    const float scale = masks[i] ? inv : 0.0f;
    if (dlogits) {
      float g[64];
      float gsum = 0.0f;
      for (int64_t j = 0; j < A; ++j) {
        const float p = expf(lp[j]);
        g[j] = ecoef * p * (lp[j] + 1.0f); /* d(-ecoef*H)/dlogp_j */
        if (j == a)
          g[j] += -g_rho * rho;            /* d(-surr)/dlogp_a */
        gsum += g[j];
      }
      for (int64_t j = 0; j < A; ++j)
        dlogits[i * A + j] = (g[j] - expf(lp[j]) * gsum) * scale; /* log-softmax backward */
    }
    if (dvalues)
      dvalues[i] = vcoef * dvl * scale;
  }
  if (loss)
    *loss = (float)(acc / count);
  free(logp);
  return RS_OK;
}

/* ---- Nature-CNN policy: src/bin/train.cc:230-265 ------------------------------------ */
typedef struct {
  int64_t hidden, actions;
  int64_t o_c1w, o_c1b, o_c2w, o_c2b, o_c3w, o_c3b, o_fcw, o_fcb, o_piw, o_pib, o_vw, o_vb, total;
} rs_layout;

static rs_layout rs_make_layout(int64_t hidden, int64_t actions) {
  rs_layout L;
  int64_t o = 0;
  L.hidden = hidden; L.actions = actions;
  L.o_c1w = o; o += 32 * 4 * 8 * 8;   L.o_c1b = o; o += 32;
  L.o_c2w = o; o += 64 * 32 * 4 * 4;  L.o_c2b = o; o += 64;
  L.o_c3w = o; o += 64 * 64 * 3 * 3;  L.o_c3b = o; o += 64;
  L.o_fcw = o; o += hidden * 3136;    L.o_fcb = o; o += hidden;
  L.o_piw = o; o += actions * hidden; L.o_pib = o; o += actions;
  L.o_vw = o;  o += hidden;           L.o_vb = o;  o += 1;
  L.total = o;
  return L;
}
int64_t rs_param_count(int64_t hidden, int64_t actions) { return rs_make_layout(hidden, actions).total; }

/* y[co,oy,ox] = relu?(b[co] + sum_{ci,ky,kx} x[ci,oy*s+ky,ox*s+kx] * w[co,ci,ky,kx]) */
static void conv_fwd(const float *x, const float *w, const float *b, float *y, int ci, int hi,
                     int wi, int co, int k, int s, int relu) {
  const int ho = (hi - k) / s + 1, wo = (wi - k) / s + 1;
  for (int o = 0; o < co; ++o)
    for (int oy = 0; oy < ho; ++oy)
      for (int ox = 0; ox < wo; ++ox) {
        float acc = b[o];
        for (int c = 0; c < ci; ++c)
          for (int ky = 0; ky < k; ++ky) {
            const float *xr = x + (c * hi + oy * s + ky) * wi + ox * s;
            const float *wr = w + ((o * ci + c) * k + ky) * k;
            for (int kx = 0; kx < k; ++kx)
              acc += xr[kx] * wr[kx];
          }
        y[(o * ho + oy) * wo + ox] = relu && acc < 0.0f ? 0.0f : acc;
      }
}
/* dy already masked by relu; accumulates dw, db; writes dx (if non-null) */
static void conv_bwd(const float *x, const float *w, const float *dy, float *dx, float *dw,
                     float *db, int ci, int hi, int wi, int co, int k, int s) {
  const int ho = (hi - k) / s + 1, wo = (wi - k) / s + 1;
  if (dx)
    memset(dx, 0, sizeof(float) * (size_t)(ci * hi * wi));
  for (int o = 0; o < co; ++o)
    for (int oy = 0; oy < ho; ++oy)
      for (int ox = 0; ox < wo; ++ox) {
        const float g = dy[(o * ho + oy) * wo + ox];
        if (g == 0.0f)
          continue;
        db[o] += g;
        for (int c = 0; c < ci; ++c)
          for (int ky = 0; ky < k; ++ky) {
            const int xi = (c * hi + oy * s + ky) * wi + ox * s;
            const int wi_ = ((o * ci + c) * k + ky) * k;
            for (int kx = 0; kx < k; ++kx) {
              dw[wi_ + kx] += g * x[xi + kx];
              if (dx)
                dx[xi + kx] += g * w[wi_ + kx];
            }
          }
      }
}

#define A0 (4 * 84 * 84)
#define A1 (32 * 20 * 20)
#define A2 (64 * 9 * 9)
#define A3 (64 * 7 * 7)

/* forward of n samples; acts (optional) receives [n, A0+A1+A2+A3+hidden] saved activations */
int rs_net_forward(const float *params, int64_t hidden, int64_t actions, const uint8_t *obs,
                   int64_t n, float *logits, float *values, float *acts) {
  const rs_layout L = rs_make_layout(hidden, actions);
  const int64_t stride = A0 + A1 + A2 + A3 + hidden;
  float *tmp = acts ? NULL : (float *)malloc(sizeof(float) * (size_t)stride);
  for (int64_t i = 0; i < n; ++i) {
    float *a0 = acts ? acts + i * stride : tmp;
    float *a1 = a0 + A0, *a2 = a1 + A1, *a3 = a2 + A2, *h = a3 + A3;
    for (int j = 0; j < A0; ++j)
      a0[j] = (float)obs[i * A0 + j] / 255.0f; /* train.cc:258-259 */
    conv_fwd(a0, params + L.o_c1w, params + L.o_c1b, a1, 4, 84, 84, 32, 8, 4, 1);
    conv_fwd(a1, params + L.o_c2w, params + L.o_c2b, a2, 32, 20, 20, 64, 4, 2, 1);
    conv_fwd(a2, params + L.o_c3w, params + L.o_c3b, a3, 64, 9, 9, 64, 3, 1, 1);
    for (int64_t o = 0; o < hidden; ++o) { /* FC 3136->hidden, NO activation (train.cc:240-241) */
      float acc = params[L.o_fcb + o];
      const float *wr = params + L.o_fcw + o * 3136;
      for (int j = 0; j < A3; ++j)
        acc += a3[j] * wr[j];
      h[o] = acc;
    }
    for (int64_t o = 0; o < actions; ++o) {
      float acc = params[L.o_pib + o];
      for (int64_t j = 0; j < hidden; ++j)
        acc += h[j] * params[L.o_piw + o * hidden + j];
      logits[i * actions + o] = acc;
    }
    float acc = params[L.o_vb];
    for (int64_t j = 0; j < hidden; ++j)
      acc += h[j] * params[L.o_vw + j];
    values[i] = acc;
  }
  free(tmp);
  return RS_OK;
}

/* backward given saved acts and d(loss)/d(logits,values); grads (flat, zeroed here) */
int rs_net_backward(const float *params, int64_t hidden, int64_t actions, const float *acts,
                    int64_t n, const float *dlogits, const float *dvalues, float *grads) {
  const rs_layout L = rs_make_layout(hidden, actions);
  const int64_t stride = A0 + A1 + A2 + A3 + hidden;
  memset(grads, 0, sizeof(float) * (size_t)L.total);
  float *dh = (float *)malloc(sizeof(float) * (size_t)hidden);
  float *d3 = (float *)malloc(sizeof(float) * A3);
  float *d2 = (float *)malloc(sizeof(float) * A2);
  float *d1 = (float *)malloc(sizeof(float) * A1);
  for (int64_t i = 0; i < n; ++i) {
    const float *a0 = acts + i * stride, *a1 = a0 + A0, *a2 = a1 + A1, *a3 = a2 + A2,
                *h = a3 + A3;
    for (int64_t j = 0; j < hidden; ++j)
      dh[j] = dvalues[i] * params[L.o_vw + j];
    grads[L.o_vb] += dvalues[i];
    for (int64_t j = 0; j < hidden; ++j)
      grads[L.o_vw + j] += dvalues[i] * h[j];
    for (int64_t o = 0; o < actions; ++o) {
      const float g = dlogits[i * actions + o];
      grads[L.o_pib + o] += g;
      for (int64_t j = 0; j < hidden; ++j) {
        grads[L.o_piw + o * hidden + j] += g * h[j];
        dh[j] += g * params[L.o_piw + o * hidden + j];
      }
    }
    memset(d3, 0, sizeof(float) * A3);
    for (int64_t o = 0; o < hidden; ++o) {
      const float g = dh[o];
      grads[L.o_fcb + o] += g;
      float *gw = grads + L.o_fcw + o * 3136;
      const float *wr = params + L.o_fcw + o * 3136;
      for (int j = 0; j < A3; ++j) {
        gw[j] += g * a3[j];
        d3[j] += g * wr[j];
      }
    }
    for (int j = 0; j < A3; ++j)
      if (a3[j] <= 0.0f) d3[j] = 0.0f; /* ReLU backward */
    conv_bwd(a2, params + L.o_c3w, d3, d2, grads + L.o_c3w, grads + L.o_c3b, 64, 9, 9, 64, 3, 1);
    for (int j = 0; j < A2; ++j)
      if (a2[j] <= 0.0f) d2[j] = 0.0f;
    conv_bwd(a1, params + L.o_c2w, d2, d1, grads + L.o_c2w, grads + L.o_c2b, 32, 20, 20, 64, 4, 2);
    for (int j = 0; j < A1; ++j)
      if (a1[j] <= 0.0f) d1[j] = 0.0f;
    conv_bwd(a0, params + L.o_c1w, d1, NULL, grads + L.o_c1w, grads + L.o_c1b, 4, 84, 84, 32, 8, 4);
  }
  free(dh); free(d3); free(d2); free(d1);
  return RS_OK;
}

/* ---- clip_grad_norm_ (src/ai/ppo/train.cc:12-46) + torch Adam (train.cc:360-362) ------
 * total = || (||g_i||_2)_i ||_2 over the 12 parameter tensors; coef = min(1, max/(total+1e-6));
 * Adam (torch form): m=b1*m+(1-b1)g; v=b2*v+(1-b2)g^2; p -= (lr/(1-b1^t)) * m/(sqrt(v)/sqrt(1-b2^t)+eps)
 * `step` is the 1-based step count AFTER increment. Returns the PRE-clip total norm. */
float rs_clip_adam(float *params, float *grads, float *m, float *v, int64_t hidden,
                   int64_t actions, int64_t step, double lr, double max_norm) {
  const rs_layout L = rs_make_layout(hidden, actions);
  const int64_t offs[13] = {L.o_c1w, L.o_c1b, L.o_c2w, L.o_c2b, L.o_c3w, L.o_c3b, L.o_fcw,
                            L.o_fcb, L.o_piw, L.o_pib, L.o_vw,  L.o_vb,  L.total};
  double tot = 0.0;
  for (int t = 0; t < 12; ++t) {
    double s = 0.0;
    for (int64_t i = offs[t]; i < offs[t + 1]; ++i)
      s += (double)grads[i] * (double)grads[i];
    const float nrm = (float)sqrt(s);
    tot += (double)nrm * (double)nrm;
  }
  const float total = (float)sqrt(tot);
  float coef = (float)(max_norm / ((double)total + 1e-6)); /* train.cc:39 */
  if (coef > 1.0f) coef = 1.0f;                             /* train.cc:40-41 */
  const double b1 = 0.9, b2 = 0.999, eps = 1e-5;
  const double bc1 = 1.0 - pow(b1, (double)step), bc2 = 1.0 - pow(b2, (double)step);
  const float step_size = (float)(lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  for (int64_t i = 0; i < L.total; ++i) {
    const float g = grads[i] * coef; /* train.cc:42-44 mul_ */
    grads[i] = g;
    m[i] = m[i] * (float)b1 + g * (float)(1.0 - b1);
    v[i] = v[i] * (float)b2 + (g * g) * (float)(1.0 - b2);
    const float denom = sqrtf(v[i]) / bc2_sqrt + (float)eps;
    params[i] = params[i] - step_size * (m[i] / denom);
  }
  return total;
}

int rs_loss(const float *logits, const float *old_log_probs, const int64_t *actions,
            const float *advantages, const float *values, const float *returns,
            const uint8_t *masks, int64_t n, int64_t A, float clip, float vcoef, float ecoef,
            float *loss, float *clipped, float *value_losses, float *entropies,
            float *total_losses, float *ratio, float *dlogits, float *dvalues) {
  return rs_loss_ex(logits, old_log_probs, actions, advantages, values, returns, masks, n, A, clip,
                    vcoef, ecoef, loss, clipped, value_losses, entropies, total_losses, ratio, dlogits,
                    dvalues, NULL);
}

/* ---- train(): src/ai/ppo/train.h:114-157 ------------------------------------------------
 * Contiguous minibatch slices (the randperm at :146 is never consumed). metrics are
 * [epochs,nmb,mb] (m_gradnorm [epochs,nmb]); `step` in/out is the Adam step counter. */
int rs_train_ex(float *params, float *m, float *v, int64_t *step, int64_t hidden, int64_t actions,
                const uint8_t *obs, const int64_t *act, const float *log_probs,
                const float *advantages, const float *returns, const uint8_t *masks, int64_t N,
                int64_t epochs, int64_t nmb, float clip, float vcoef, float ecoef, float max_norm,
                double lr, float *m_loss, float *m_clipped, float *m_value, float *m_entropy,
                float *m_ratio, float *m_total, float *m_adv, float *m_ret, uint8_t *m_masks,
                float *m_gradnorm, const float *old_values /* NULL = reference value loss */) {
  if (nmb <= 0 || N % nmb != 0)
    return RS_RUNTIME_ERROR; /* train.h:140-143 */
  const int64_t mb = N / nmb, P = rs_param_count(hidden, actions);
  const int64_t stride = A0 + A1 + A2 + A3 + hidden;
  float *acts = (float *)malloc(sizeof(float) * (size_t)(mb * stride));
  float *logits = (float *)malloc(sizeof(float) * (size_t)(mb * actions));
  float *vals = (float *)malloc(sizeof(float) * (size_t)mb);
  float *dlog = (float *)malloc(sizeof(float) * (size_t)(mb * actions));
  float *dval = (float *)malloc(sizeof(float) * (size_t)mb);
  float *grads = (float *)malloc(sizeof(float) * (size_t)P);
  for (int64_t ep = 0; ep < epochs; ++ep)
    for (int64_t k = 0; k < nmb; ++k) {
      const int64_t s = k * mb, o = (ep * nmb + k) * mb;
      float loss;
      rs_net_forward(params, hidden, actions, obs + s * A0, mb, logits, vals, acts);
      rs_loss_ex(logits, log_probs + s * actions, act + s, advantages + s, vals, returns + s,
                 masks + s, mb, actions, clip, vcoef, ecoef, &loss, m_clipped ? m_clipped + o : NULL,
                 m_value ? m_value + o : NULL, m_entropy ? m_entropy + o : NULL,
                 m_total ? m_total + o : NULL, m_ratio ? m_ratio + o : NULL, dlog, dval,
                 old_values ? old_values + s : NULL);
      rs_net_backward(params, hidden, actions, acts, mb, dlog, dval, grads);
      *step += 1;
      const float tn = rs_clip_adam(params, grads, m, v, hidden, actions, *step, lr, max_norm);
      for (int64_t i = 0; i < mb; ++i) { /* Metrics::set, train.h:93-108 */
        if (m_loss) m_loss[o + i] = loss;
        if (m_adv) m_adv[o + i] = advantages[s + i];
        if (m_ret) m_ret[o + i] = returns[s + i];
        if (m_masks) m_masks[o + i] = masks[s + i];
      }
      if (m_gradnorm) m_gradnorm[ep * nmb + k] = tn;
    }
  free(acts); free(logits); free(vals); free(dlog); free(dval); free(grads);
  return RS_OK;
}

int rs_train(float *params, float *m, float *v, int64_t *step, int64_t hidden, int64_t actions,
             const uint8_t *obs, const int64_t *act, const float *log_probs,
             const float *advantages, const float *returns, const uint8_t *masks, int64_t N,
             int64_t epochs, int64_t nmb, float clip, float vcoef, float ecoef, float max_norm,
             double lr, float *m_loss, float *m_clipped, float *m_value, float *m_entropy,
             float *m_ratio, float *m_total, float *m_adv, float *m_ret, uint8_t *m_masks,
             float *m_gradnorm) {
  return rs_train_ex(params, m, v, step, hidden, actions, obs, act, log_probs, advantages, returns,
                     masks, N, epochs, nmb, clip, vcoef, ecoef, max_norm, lr, m_loss, m_clipped,
                     m_value, m_entropy, m_ratio, m_total, m_adv, m_ret, m_masks, m_gradnorm, NULL);
}
