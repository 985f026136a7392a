// Host env feed: what sits between the emulator workers and the rollout buffers, as batched native code.
//
// The reference assembles every observation in Python, per env and per step: OCAtari.ns_state chains the [x, y] (or
// [0, 0] when invisible, ocatari/ram/game_objects.py:184-188) of all object slots into a list
// (ocatari/core.py:444-448), appends it to a deque(maxlen=buffer_window_size) (core.py:152-153, 274-283; a reset fills
// all entries with the initial state, core.py:241-244) and returns np.array(deque) (core.py:213-214, oldest frame
// first); SubprocVecEnv stacks the N results, VecNormalize(norm_obs=False, norm_reward=True) rescales the rewards by
// the running std of the discounted returns (cleanrl/ppo_atari_oc.py:411-414; stable-baselines3==2.0.0,
// common/vec_env/vec_normalize.py + common/running_mean_std.py -- not vendored in the reference tree, algorithm
// restated in oracle/ref_feed.py), and the script builds three pageable tensors per step (ppo_atari_oc.py:512-514).
// Here one call per step takes the object slots of ALL envs and writes the (N, frames, 2 * slots) fp32 block straight
// into the caller's buffer (the rollout collector's pinned staging memory), and one call normalises the rewards.
// Pure host code: no CUDA calls, no threads.
#include "../../include/ocppo.h"

#include <math.h>
#include <new>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <vector>

namespace ocppo { void set_error(const char *fmt, ...); }

struct ocppo_feed {
    int32_t num_envs, frames, slots, dim;       // dim = 2 * slots
    std::vector<float> ring;                    // [N][frames][dim] frame ring per env
    std::vector<int32_t> head;                  // index of the OLDEST frame of env n
    std::vector<uint8_t> primed;                // env has been filled at least once
    // VecNormalize(norm_reward=True) state
    double gamma, epsilon, clip_reward;
    std::vector<double> returns;                // [N] discounted returns
    double rms_mean, rms_var, rms_count;        // RunningMeanStd(shape=()) of the returns
};

namespace {

#define FEED_CHECK(cond, ...)                \
    do {                                     \
        if (!(cond)) {                       \
            ::ocppo::set_error(__VA_ARGS__); \
            return OCPPO_EINVAL;             \
        }                                    \
    } while (0)

// numpy's pairwise summation (numpy/core/src/umath/loops_utils.h.src, DOUBLE_pairwise_sum, contiguous case): the
// reduction order np.mean / np.var use, so the running statistics equal the reference's bit for bit
double pairwise_sum(const double *a, int64_t n) {
    if (n < 8) {
        double res = 0.;
        for (int64_t i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int64_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    int64_t n2 = n / 2;
    n2 -= n2 % 8;
    return pairwise_sum(a, n2) + pairwise_sum(a + n2, n - n2);
}

}  // namespace

extern "C" {

int ocppo_feed_create(int32_t num_envs, int32_t frames, int32_t slots, double gamma, double epsilon, double clip_reward,
                      ocppo_feed **out) {
    FEED_CHECK(out != nullptr, "feed_create: null argument");
    FEED_CHECK(num_envs >= 1 && frames >= 1 && slots >= 1, "feed_create: num_envs, frames and slots must be >= 1");
    ocppo_feed *f = new (std::nothrow) ocppo_feed();
    FEED_CHECK(f != nullptr, "feed_create: out of host memory");
    f->num_envs = num_envs; f->frames = frames; f->slots = slots; f->dim = 2 * slots;
    f->ring.assign((size_t)num_envs * frames * f->dim, 0.f);
    f->head.assign(num_envs, 0);
    f->primed.assign(num_envs, 0);
    f->gamma = gamma; f->epsilon = epsilon; f->clip_reward = clip_reward;
    f->returns.assign(num_envs, 0.0);
    f->rms_mean = 0.0; f->rms_var = 1.0; f->rms_count = 1e-4;       // RunningMeanStd(epsilon=1e-4, shape=())
    *out = f;
    return OCPPO_OK;
}

int ocppo_feed_destroy(ocppo_feed *f) {
    delete f;
    return OCPPO_OK;
}

int ocppo_feed_push(ocppo_feed *f, const int32_t *xy, const uint8_t *visible, const uint8_t *refill, float *obs_out) {
    FEED_CHECK(f != nullptr && xy != nullptr && visible != nullptr && obs_out != nullptr, "feed_push: null argument");
    const int F = f->frames, S = f->slots, D = f->dim;
    for (int32_t n = 0; n < f->num_envs; ++n) {
        float *ring = f->ring.data() + (size_t)n * F * D;
        const int32_t *p = xy + (size_t)n * S * 2;
        const uint8_t *v = visible + (size_t)n * S;
        const bool fill = (refill != nullptr && refill[n]) || !f->primed[n];
        // the new neurosymbolic state: [x, y] per slot, [0, 0] when the slot's object is not visible
        float *dst = ring + (size_t)f->head[n] * D;           // overwrites the oldest frame (deque maxlen semantics)
        for (int s = 0; s < S; ++s) {
            dst[2 * s] = v[s] ? (float)p[2 * s] : 0.f;
            dst[2 * s + 1] = v[s] ? (float)p[2 * s + 1] : 0.f;
        }
        if (fill) {                                           // reset: every entry of the stack is the current state
            for (int k = 0; k < F; ++k)
                if (k != f->head[n]) memcpy(ring + (size_t)k * D, dst, (size_t)D * sizeof(float));
            f->head[n] = 0;
            f->primed[n] = 1;
        } else {
            f->head[n] = (f->head[n] + 1) % F;                // the frame after the newest is now the oldest
        }
        float *o = obs_out + (size_t)n * F * D;               // np.array(deque): oldest frame first
        const int h = f->head[n];
        memcpy(o, ring + (size_t)h * D, (size_t)(F - h) * D * sizeof(float));
        if (h > 0) memcpy(o + (size_t)(F - h) * D, ring, (size_t)h * D * sizeof(float));
    }
    return OCPPO_OK;
}

int ocppo_feed_normalize_reward(ocppo_feed *f, const double *reward, const uint8_t *done, float *reward_out, float *done_out) {
    FEED_CHECK(f != nullptr && reward != nullptr && done != nullptr && reward_out != nullptr, "feed_normalize_reward: null argument");
    const int64_t N = f->num_envs;
    // VecNormalize._update_reward: returns = returns * gamma + reward; ret_rms.update(returns)
    for (int64_t n = 0; n < N; ++n) f->returns[n] = f->returns[n] * f->gamma + reward[n];
    // RunningMeanStd.update: batch_mean = np.mean(x), batch_var = np.var(x) (numpy's reduction order), then
    // update_from_moments
    const double batch_mean = pairwise_sum(f->returns.data(), N) / (double)N;
    std::vector<double> dev2(N);
    for (int64_t n = 0; n < N; ++n) {
        const double d = f->returns[n] - batch_mean;
        dev2[n] = d * d;
    }
    const double batch_var = pairwise_sum(dev2.data(), N) / (double)N;
    const double batch_count = (double)N;
    const double delta = batch_mean - f->rms_mean;
    const double tot_count = f->rms_count + batch_count;
    const double new_mean = f->rms_mean + delta * batch_count / tot_count;
    const double m_a = f->rms_var * f->rms_count;
    const double m_b = batch_var * batch_count;
    const double m_2 = m_a + m_b + (delta * delta) * f->rms_count * batch_count / (f->rms_count + batch_count);
    f->rms_var = m_2 / (f->rms_count + batch_count);
    f->rms_mean = new_mean;
    f->rms_count = batch_count + f->rms_count;
    // normalize_reward: clip(reward / sqrt(var + epsilon), -clip, clip); then the script's float32 cast (:512)
    const double denom = sqrt(f->rms_var + f->epsilon);
    for (int64_t n = 0; n < N; ++n) {
        double r = reward[n] / denom;
        r = r < -f->clip_reward ? -f->clip_reward : (r > f->clip_reward ? f->clip_reward : r);
        reward_out[n] = (float)r;
        if (done_out != nullptr) done_out[n] = done[n] ? 1.f : 0.f;     // next_done as float32 (:514)
        if (done[n]) f->returns[n] = 0.0;                               // self.returns[dones] = 0
    }
    return OCPPO_OK;
}

int ocppo_feed_state(const ocppo_feed *f, double *mean_var_count3, double *returns) {
    FEED_CHECK(f != nullptr, "feed_state: null handle");
    if (mean_var_count3 != nullptr) { mean_var_count3[0] = f->rms_mean; mean_var_count3[1] = f->rms_var; mean_var_count3[2] = f->rms_count; }
    if (returns != nullptr) memcpy(returns, f->returns.data(), (size_t)f->num_envs * sizeof(double));
    return OCPPO_OK;
}

}  // extern "C"
