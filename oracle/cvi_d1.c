/*
 * Oracle, state dimension 1, in C: one CVI-DP site update (cvi_dp_trainer.py:71-77) per sequence, the
 * sequences spread over OpenMP threads.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): used by tests/ (against the NumPy oracle it restates,
 * oracle/scalar_d1.py) and by bench.py as the timed CPU baseline.  Nothing here is shipped or called by the
 * product path.
 *
 * Statement for statement the same operations, in the same order, as oracle/scalar_d1.py, which follows
 *   ssm_to_naturals           markovflow/ssm_gaussian_transformations.py:182-253
 *   naturals_to_ssm_params    markovflow/ssm_gaussian_transformations.py:333-511 (cholesky_band,
 *                             inverse_from_cholesky_band of banded_matrices 0.0.6)
 *   marginals                 markovflow/state_space_model.py:231-262, 431-483
 *   KL + gradient             markovflow/sde/sde_utils.py:473-547 (through :93-178 and :262-359), mvnquad H = 20
 *   KL_q_p / classic_elbo     markovflow/models/variational_cvi_sde.py:339-352, 446-486
 *   site updates              markovflow/models/variational_cvi_sde.py:279-317
 * Build: gcc -O2 -fopenmp -shared -fPIC (no -ffast-math, no FMA contraction), see oracle/build_c.sh.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { DRIFT_OU = 0, DRIFT_DW = 1, DRIFT_BENES = 2, DRIFT_SINE = 3, DRIFT_SQRT = 4 };

/* sde.py drifts (D = 1) and their derivative */
static inline void drift_and_grad(int id, const double* p, double x, double* f, double* fp) {
    switch (id) {
    case DRIFT_OU: *f = -p[0] * x; *fp = -p[0]; break;
    case DRIFT_DW: *f = p[0] * x * (p[1] - x * x); *fp = p[0] * (p[1] - 3.0 * (x * x)); break;
    case DRIFT_BENES: { const double t = tanh(x); *f = p[0] * t; *fp = p[0] * (1.0 - t * t); break; }
    case DRIFT_SINE: *f = sin(x - p[0]); *fp = cos(x - p[0]); break;
    default: { const double r = sqrt(p[0] * fabs(x)); *f = r; *fp = 0.5 / r * p[0] * ((x > 0) - (x < 0)); break; }
    }
}

typedef struct {
    int64_t n1, n_obs;               /* grid points N+1, observations */
    int drift_id;
    const double* params;
    double q, dt, noise_std, p0_mu, p0_cov;
    const double* time_grid;         /* [n1] */
    const double *gh_x, *gh_w;       /* Gauss-Hermite nodes / weights, H = 20 */
    int H;
} Problem;

typedef struct {                     /* per-thread scratch, all [n1] */
    double *th1, *thd, *ths, *l, *ls, *Sd, *Ss, *A, *b, *cQ, *m, *S, *s1, *s2, *g1, *g2, *g3, *y, *tmp;
    double mu0, cP0;
} Work;

static int work_alloc(Work* w, int64_t n1) {
    double** f[] = {&w->th1, &w->thd, &w->ths, &w->l, &w->ls, &w->Sd, &w->Ss, &w->A, &w->b, &w->cQ,
                    &w->m, &w->S, &w->s1, &w->s2, &w->g1, &w->g2, &w->g3, &w->y, &w->tmp};
    for (size_t i = 0; i < sizeof(f) / sizeof(f[0]); ++i) {
        *f[i] = (double*)malloc(sizeof(double) * (size_t)n1);
        if (!*f[i]) return -1;
    }
    return 0;
}
static void work_free(Work* w) {
    double* f[] = {w->th1, w->thd, w->ths, w->l, w->ls, w->Sd, w->Ss, w->A, w->b, w->cQ,
                   w->m, w->S, w->s1, w->s2, w->g1, w->g2, w->g3, w->y, w->tmp};
    for (size_t i = 0; i < sizeof(f) / sizeof(f[0]); ++i) free(f[i]);
}

/* cholesky_band (block_tri_diag.py:440) */
static void chol_fwd(const double* Kd, const double* Ks, double* l, double* ls, int64_t n) {
    l[0] = sqrt(Kd[0]);
    for (int64_t k = 0; k < n - 1; ++k) {
        ls[k] = Ks[k] / l[k];
        l[k + 1] = sqrt(Kd[k + 1] - ls[k] * ls[k]);
    }
}

/* inverse_from_cholesky_band (ssm_gaussian_transformations.py:443-445) */
static void inv_band_bwd(const double* l, const double* ls, double* Sd, double* Ss, int64_t n) {
    double U = 1.0 / l[n - 1];
    Sd[n - 1] = U * U;
    for (int64_t k = n - 2; k >= 0; --k) {
        U = 1.0 / l[k];
        const double G = ls[k] * U;
        Ss[k] = -Sd[k + 1] * G;
        Sd[k] = U * U - Ss[k] * G;
    }
}

/* dist_q: prior naturals (ssm_to_naturals) + girsanov sites + scattered data sites -> q as an SSM
 * (naturals_to_ssm_params); the result goes to w->A, w->b, w->cQ, w->mu0, w->cP0 */
static void dist_q(const Problem* P, Work* w, const double* pA, const double* pb, const double* pcQ, double pmu0,
                   double pcP0, const double* g1, const double* gd, const double* gs, const double* d1,
                   const double* d2, const int64_t* obs_idx) {
    const int64_t n = P->n1;
    /* ssm_to_naturals */
    for (int64_t k = 0; k < n; ++k) {
        const double c = (k == 0) ? pcP0 : pcQ[k - 1];
        const double o = (k == 0) ? pmu0 : pb[k - 1];
        w->tmp[k] = (o / c) / c;
    }
    for (int64_t k = 0; k < n; ++k) {
        const double c = (k == 0) ? pcP0 : pcQ[k - 1];
        const double prec = (1.0 / c) / c;
        if (k < n - 1) {
            const double LinvA = pA[k] / pcQ[k];
            w->ths[k] = LinvA / pcQ[k];
            w->th1[k] = w->tmp[k] - pA[k] * w->tmp[k + 1];
            w->thd[k] = -0.5 * (prec + LinvA * LinvA);
        } else {
            w->th1[k] = w->tmp[k];
            w->thd[k] = -0.5 * (prec + 0.0);
        }
    }
    /* + girsanov sites + scatter(data sites)   (variational_cvi_sde.py:161-174) */
    memset(w->s1, 0, sizeof(double) * (size_t)n);
    memset(w->s2, 0, sizeof(double) * (size_t)n);
    for (int64_t j = 0; j < P->n_obs; ++j) { w->s1[obs_idx[j]] += d1[j]; w->s2[obs_idx[j]] += d2[j]; }
    for (int64_t k = 0; k < n; ++k) {
        w->th1[k] = w->th1[k] + g1[k] + w->s1[k];
        w->thd[k] = w->thd[k] + gd[k] + w->s2[k];
        if (k < n - 1) w->ths[k] = w->ths[k] + gs[k];
    }
    /* naturals_to_ssm_params */
    double* Kd = w->tmp;       /* -2 theta_d */
    double* Ks = w->y;         /* -theta_s   */
    for (int64_t k = 0; k < n; ++k) { Kd[k] = -2.0 * w->thd[k]; if (k < n - 1) Ks[k] = -w->ths[k]; }
    chol_fwd(Kd, Ks, w->l, w->ls, n);
    inv_band_bwd(w->l, w->ls, w->Sd, w->Ss, n);
    for (int64_t k = 0; k < n - 1; ++k) w->A[k] = w->Ss[k] / w->Sd[k];
    /* conditional precisions -> covariances, their Cholesky factors, offsets */
    double* cov = w->S;        /* reused: overwritten by the marginals later */
    double* chols = w->m;
    for (int64_t k = 0; k < n; ++k) {
        const double cp = (k < n - 1) ? Kd[k] + w->A[k] * Ks[k] : Kd[k];
        const double ccp = sqrt(cp);
        cov[k] = (1.0 / ccp) / ccp;
        chols[k] = sqrt(cov[k]);
    }
    double yk = w->th1[n - 1];
    double off_last = cov[n - 1] * yk;
    /* offsets b[k-1] = cov[k] * y[k], y[k] = th1[k] + A[k] y[k+1] */
    if (n >= 2) w->b[n - 2] = off_last;
    for (int64_t k = n - 2; k >= 0; --k) {
        yk = w->th1[k] + w->A[k] * yk;
        const double off = cov[k] * yk;
        if (k >= 1) w->b[k - 1] = off; else w->mu0 = off;
    }
    if (n == 1) w->mu0 = off_last;
    w->cP0 = chols[0];
    for (int64_t k = 1; k < n; ++k) w->cQ[k - 1] = chols[k];
}

/* marginals of the SSM in w (state_space_model.py:231-262, 431-483) -> m, S */
static void ssm_marginals(Work* w, int64_t n, double* m, double* S) {
    m[0] = w->mu0;
    for (int64_t k = 0; k < n - 1; ++k) m[k + 1] = w->b[k] + w->A[k] * m[k];
    double* diag = w->tmp;
    double* sub = w->y;
    for (int64_t k = 0; k < n; ++k) {
        const double c = (k == 0) ? w->cP0 : w->cQ[k - 1];
        const double cinv = (1.0 / c) / c;
        if (k < n - 1) {
            const double inv_q_a = (w->A[k] / w->cQ[k]) / w->cQ[k];
            diag[k] = cinv + w->A[k] * inv_q_a;
            sub[k] = -inv_q_a;
        } else {
            diag[k] = cinv + 0.0;
        }
    }
    chol_fwd(diag, sub, w->l, w->ls, n);
    inv_band_bwd(w->l, w->ls, S, w->Ss, n);
}

/* sde_utils.py:473-547 for D = 1: KL[q || p_SDE] (not returned) and d KL / d(eta1, eta_d, eta_s) -> w->g1, g2, g3.
 * q is the SSM in w with marginals (m, S).  Array passes in the order of scalar_d1.kl_and_grads. */
static void kl_grads(const Problem* P, Work* w, const double* m, const double* S) {
    const int64_t n = P->n1, N = n - 1;
    const double sqrt2 = sqrt(2.0);
    double* g_m = w->g1;
    double* g_S = w->g2;
    double* g_C = w->g3;
    double* Sc = w->s1;     /* ed - e1^2 */
    double* LS = w->s2;
    double* g_b = w->th1;   /* scratch arrays reused: the naturals are dead once q is an SSM */
    double* g_A = w->thd;
    double* sXb = w->ths;
    double* LSb = w->l;
    double* gQc = w->ls;
    double* Aq = w->Sd;     /* A recomputed from the expectation parameters */
    for (int64_t k = 0; k < n; ++k) {
        const double ed = S[k] + m[k] * m[k];
        Sc[k] = ed - m[k] * m[k];
        LS[k] = sqrt(Sc[k]);
        g_m[k] = 0.0;
        g_S[k] = 0.0;
    }
    for (int64_t k = 0; k < N; ++k) {
        const double es = w->A[k] * S[k] + m[k + 1] * m[k];
        const double Cs = es - m[k] * m[k + 1];
        const double A = (Cs / LS[k]) / LS[k];
        const double b = m[k + 1] - A * m[k];
        const double Qc = Sc[k + 1] - A * (Sc[k] * A);
        const double LQ = sqrt(Qc);
        const double Qq = LQ * LQ;
        const double Qp = (P->time_grid[k + 1] - P->time_grid[k]) * P->q;
        const double cQp = sqrt(Qp);
        const double Qpinv = (1.0 / cQp) / cQp;
        double s_rb = 0.0, s_rbX = 0.0, s_Xb = 0.0, s_Xbx = 0.0;
        for (int p = 0; p < P->H; ++p) {
            const double X = sqrt2 * (LS[k] * P->gh_x[p]) + m[k];
            double f, fp;
            drift_and_grad(P->drift_id, P->params, X, &f, &fp);
            const double d = (X + P->dt * f) - (A * X + b);
            const double r = Qpinv * d;
            const double rb = P->gh_w[p] * r;
            const double Xb = ((1.0 + P->dt * fp) - A) * rb;
            s_rb += rb;
            s_rbX += rb * X;
            s_Xb += Xb;
            s_Xbx += Xb * P->gh_x[p];
        }
        Aq[k] = A;
        g_b[k] = -s_rb;
        g_A[k] = -s_rbX;
        sXb[k] = s_Xb;
        LSb[k] = sqrt2 * s_Xbx;
        gQc[k] = 0.5 * (-1.0 / Qq + Qpinv);
    }
    for (int64_t k = 0; k < N; ++k) g_m[k] += sXb[k];
    for (int64_t k = 0; k < N; ++k) g_S[k] += LSb[k] / (2.0 * LS[k]);
    for (int64_t k = 0; k < N; ++k) g_m[k + 1] += g_b[k];
    for (int64_t k = 0; k < N; ++k) g_A[k] += -g_b[k] * m[k];
    for (int64_t k = 0; k < N; ++k) g_m[k] += -Aq[k] * g_b[k];
    for (int64_t k = 0; k < N; ++k) g_S[k + 1] += gQc[k];
    for (int64_t k = 0; k < N; ++k) g_A[k] += -2.0 * gQc[k] * Aq[k] * Sc[k];
    for (int64_t k = 0; k < N; ++k) g_S[k] += -Aq[k] * gQc[k] * Aq[k];
    for (int64_t k = 0; k < N; ++k) g_C[k] = (g_A[k] / LS[k]) / LS[k];
    for (int64_t k = 0; k < N; ++k) g_S[k] += -g_C[k] * Aq[k];
    {
        const double P0 = LS[0] * LS[0];
        g_m[0] += (m[0] - P->p0_mu) / P->p0_cov;
        g_S[0] += 0.5 * (1.0 / P->p0_cov - 1.0 / P0);
    }
    for (int64_t k = 0; k < N; ++k) g_m[k] += -g_C[k] * m[k + 1];
    for (int64_t k = 0; k < N; ++k) g_m[k + 1] += -g_C[k] * m[k];
    for (int64_t k = 0; k < n; ++k) g_m[k] += -2.0 * g_S[k] * m[k];
}

/* variational_cvi_sde.py:446-486 for D = 1 */
static double kl_q_p(const Problem* P, const Work* w, const double* m, const double* S, double p_mu0, double p_cP0) {
    const int64_t N = P->n1 - 1;
    const double sqrt2 = sqrt(2.0);
    const double Qp = P->dt * P->q;
    const double cQp = sqrt(Qp);
    const double Qpinv = (1.0 / cQp) / cQp;
    double acc = 0.0;
    for (int64_t k = 0; k < N; ++k) {
        const double Qq = w->cQ[k] * w->cQ[k];
        const double Cterm = -(2.0 * log(sqrt(Qq)) - 2.0 * log(cQp)) - 1.0 + Qpinv * Qq;
        const double sS = sqrt(S[k]);
        double quad = 0.0;
        for (int p = 0; p < P->H; ++p) {
            const double X = sqrt2 * (sS * P->gh_x[p]) + m[k];
            double f, fp;
            drift_and_grad(P->drift_id, P->params, X, &f, &fp);
            const double d = (X + P->dt * f) - (w->A[k] * X + w->b[k]);
            quad += P->gh_w[p] * (d * Qpinv * d);
        }
        acc += quad + Cterm;
    }
    const double P0 = w->cP0 * w->cP0, Pp = p_cP0 * p_cP0;
    const double kl0 = 0.5 * (P0 / Pp + (p_mu0 - w->mu0) * (p_mu0 - w->mu0) / Pp - 1.0 + log(Pp) - log(P0));
    return 0.5 * acc + kl0;
}

/*
 * B independent site updates (update_data_sites, update_girsanov_sites, classic_elbo).
 * Layouts as the CUDA library for D = 1: prior pA, pb, pcQ [B, N], pmu0, pcP0 [B]; sites g1, gd [B, N+1],
 * gs [B, N]; data sites d1, d2 [B, n_obs]; obs_idx [B, n_obs] (sorted grid indices), obs_y [B, n_obs];
 * fx_mus, fx_covs [B, N+1]; elbo [B].  gh_w already holds w_i / sqrt(pi).  Returns 0, or -1 if out of memory.
 */
int cvi_d1_site_update(int drift_id, const double* params, double q, const double* time_grid, int64_t n1, int64_t B,
                       int64_t n_obs, const int64_t* obs_idx, const double* obs_y, double noise_std, double p0_mu,
                       double p0_cov, const double* pA, const double* pb, const double* pcQ, const double* pmu0,
                       const double* pcP0, double* g1, double* gd, double* gs, double* d1, double* d2,
                       double* fx_mus, double* fx_covs, double lr_data, double lr_girsanov, const double* gh_x,
                       const double* gh_w, int H, double* elbo, int nthreads) {
    Problem P;
    P.n1 = n1; P.n_obs = n_obs; P.drift_id = drift_id; P.params = params; P.q = q;
    P.dt = time_grid[1] - time_grid[0];
    P.noise_std = noise_std; P.p0_mu = p0_mu; P.p0_cov = p0_cov; P.time_grid = time_grid;
    P.gh_x = gh_x; P.gh_w = gh_w; P.H = H;
    const int64_t N = n1 - 1;
    int failed = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        Work w;
        memset(&w, 0, sizeof(w));
        const int bad = work_alloc(&w, n1);
        if (bad) {
#pragma omp atomic write
            failed = 1;
        }
#pragma omp for schedule(dynamic, 1)
        for (int64_t s = 0; s < B; ++s) {
            if (bad) continue;
            const double *A_p = pA + s * N, *b_p = pb + s * N, *cQ_p = pcQ + s * N;
            double *G1 = g1 + s * n1, *Gd = gd + s * n1, *Gs = gs + s * N;
            double *D1 = d1 + s * n_obs, *D2 = d2 + s * n_obs;
            const int64_t* idx = obs_idx + s * n_obs;
            const double* y = obs_y + s * n_obs;
            double *mu = fx_mus + s * n1, *cov = fx_covs + s * n1;
            const double inv = (1.0 / noise_std) / noise_std;
            /* update_data_sites  variational_cvi_sde.py:301-317 */
            for (int64_t j = 0; j < n_obs; ++j) {
                const double mj = mu[idx[j]];
                const double g_mu = inv * (y[j] - mj);
                const double g_S = -0.5 * inv;
                const double gj = g_mu - 2.0 * g_S * mj;
                D1[j] = (1 - lr_data) * D1[j] + lr_data * gj;
                D2[j] = (1 - lr_data) * D2[j] + lr_data * g_S;
            }
            dist_q(&P, &w, A_p, b_p, cQ_p, pmu0[s], pcP0[s], G1, Gd, Gs, D1, D2, idx);
            ssm_marginals(&w, n1, mu, cov);
            /* update_girsanov_sites  :279-299 */
            dist_q(&P, &w, A_p, b_p, cQ_p, pmu0[s], pcP0[s], G1, Gd, Gs, D1, D2, idx);
            ssm_marginals(&w, n1, w.m, w.S);
            kl_grads(&P, &w, w.m, w.S);
            memset(w.tmp, 0, sizeof(double) * (size_t)n1);
            memset(w.y, 0, sizeof(double) * (size_t)n1);
            for (int64_t j = 0; j < n_obs; ++j) { w.tmp[idx[j]] += D1[j]; w.y[idx[j]] += D2[j]; }
            for (int64_t k = 0; k < n1; ++k) {
                G1[k] = G1[k] + lr_girsanov * (w.tmp[k] - w.g1[k]);
                Gd[k] = Gd[k] + lr_girsanov * (w.y[k] - w.g2[k]);
                if (k < N) Gs[k] = Gs[k] - lr_girsanov * w.g3[k];
            }
            dist_q(&P, &w, A_p, b_p, cQ_p, pmu0[s], pcP0[s], G1, Gd, Gs, D1, D2, idx);
            ssm_marginals(&w, n1, mu, cov);
            /* classic_elbo  :339-352 */
            dist_q(&P, &w, A_p, b_p, cQ_p, pmu0[s], pcP0[s], G1, Gd, Gs, D1, D2, idx);
            ssm_marginals(&w, n1, w.m, w.S);
            double ve = 0.0;
            for (int64_t j = 0; j < n_obs; ++j) {
                const double z = (y[j] - w.m[idx[j]]) / noise_std;
                ve += -0.5 * inv * w.S[idx[j]] + (-0.5 * z * z - log(fabs(noise_std)) - 0.5 * log(2.0 * M_PI));
            }
            elbo[s] = ve - kl_q_p(&P, &w, w.m, w.S, pmu0[s], pcP0[s]);
        }
        work_free(&w);
    }
    return failed ? -1 : 0;
}

int cvi_d1_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
