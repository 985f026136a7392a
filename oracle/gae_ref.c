/* Plain-C restatement of the reference GAE loop.  TEST INFRASTRUCTURE ONLY
 * (see oracle/__init__.py): used as the checker in tests/, smoke() and as the
 * scalar "port" CPU baseline in bench.py.  Never linked into the product.
 *
 * Follows /root/reference/cleanrl/ppo_atari_oc.py:533-547 (identical math at
 * cleanrl/ppo_atari_multigpu.py:288-301):
 *     nextnonterminal = 1 - done_{t+1}          (next_done at t = T-1)
 *     delta = r_t + gamma * V_{t+1} * nnt - V_t
 *     A_t   = delta + gamma * lambda * nnt * A_{t+1}
 *     R_t   = A_t + V_t
 * with torch's fp32 evaluation order: python scalars
 * are rounded to fp32 per op, gamma*lambda is multiplied in double first, no
 * FMA contraction (build with -ffp-contract=off).
 *
 * Parity: pinned against the torch loop restated in oracle/ref_ppo.py:ref_gae
 * (bit-exact, tests/test_oracle_cpu.py); the reference ships no GAE vectors.
 */
#include <stddef.h>
#include <stdint.h>

int ocppo_oracle_gae_f32(const float *rewards, const float *values, const float *dones,
                         const float *next_value, const float *next_done,
                         int64_t T, int64_t N, double gamma, double gae_lambda,
                         float *advantages, float *returns)
{
    if (T < 0 || N < 0) return -1;
    const float g = (float)gamma;
    const float c = (float)(gamma * gae_lambda);
    for (int64_t n = 0; n < N; ++n) {
        float last = 0.0f;
        float nv = next_value[n];
        float nnt = 1.0f - next_done[n];
        for (int64_t t = T - 1; t >= 0; --t) {
            const size_t i = (size_t)t * (size_t)N + (size_t)n;
            if (t != T - 1) {
                nv = values[i + (size_t)N];
                nnt = 1.0f - dones[i + (size_t)N];
            }
            volatile float gv = g * nv;           /* (gamma * nextvalues)        */
            volatile float gvn = gv * nnt;        /*   * nextnonterminal         */
            volatile float rp = rewards[i] + gvn; /* rewards[t] + ...            */
            const float delta = rp - values[i];   /*   - values[t]               */
            volatile float cn = c * nnt;          /* (gamma*lambda) * nnt        */
            volatile float cl = cn * last;        /*   * lastgaelam              */
            last = delta + cl;
            advantages[i] = last;
            returns[i] = last + values[i];
        }
    }
    return 0;
}
