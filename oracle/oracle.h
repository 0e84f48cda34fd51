/* oracle/oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's (smithlabcode/epievo @ 6e0525a)
 * algorithm for the MCMC / sufficient-statistics hot path.  Every function
 * cites the reference file:line it follows.  Only tests/, the smoke() check of
 * __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (epievo_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_ref.py checks every function
 * below against oracle/_ref/libepievo_ref.so (the unmodified reference sources
 * compiled by oracle/Makefile) -- bit-exact, including whole run_mcmc chains
 * driven by the same mt19937 seed -- and tests/golden/ holds vectors generated
 * from that reference build (tests/golden/make_golden.py) for machines where
 * /root/reference is absent.
 *
 * Flat "node-major CSR" path format (identical to include/epievo_b200.h):
 *   init_state uint8 [n_nodes*n_sites], row_off uint64 [n_nodes*n_sites+1],
 *   jumps double[], tot_time double[n_nodes]; row = node*n_sites + site;
 *   node 0 (root) rows are empty.
 */
#ifndef EPIEVO_ORACLE_H
#define EPIEVO_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* a3 */
void orc_trans_prob_mat(double r0, double r1, double t, double *P);
double orc_trans_prob(double r0, double r1, double t, int a, int b);
void orc_expectations(double r0, double r1, double T,
                      double *J0, double *J1, double *D0, double *D1);
/* a2 */
uint64_t orc_segment_info(const double *rates, int l_init, uint64_t nl,
                          const double *lj, int r_init, uint64_t nr,
                          const double *rj, double tot_time,
                          double *rate0, double *rate1, uint64_t *trip0,
                          uint64_t *trip1, double *len);
/* a12 */
void orc_add_suff_stats(int l_init, uint64_t nl, const double *lj,
                        int m_init, uint64_t nm, const double *mj,
                        int r_init, uint64_t nr, const double *rj,
                        double tot_time, double *J, double *D);

/* mutable path container (one global instance, like ref_driver.cpp) */
void orc_load_paths(uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
                    const uint64_t *row_off, const double *jumps,
                    const double *tot_time);
uint64_t orc_total_jumps(void);
void orc_store_paths(uint8_t *init_state, uint64_t *row_off, double *jumps,
                     double *tot_time);
void orc_free_paths(void);

/* a13, a8/a10, a2+a4+a7 */
void orc_get_suff_stats(double *J, double *D);
void orc_tri_llh(const double *rates, const double *T, double *tri_llh);
uint64_t orc_site_debug(uint64_t n_nodes, const uint64_t *subtree_sizes,
                        const uint64_t *parent_ids, const double *branches,
                        const double *rates, const double *T, uint64_t site,
                        uint64_t cap, uint64_t *seg_off, double *rate0,
                        double *rate1, uint64_t *trip0, uint64_t *trip1,
                        double *len, double *p, double *q, double *root_p0,
                        double *proposal_logp);

/* a11: sequential sampler driven by a bit-compatible mt19937 */
void orc_seed(uint64_t seed);
double orc_run_mcmc(uint64_t n_nodes, const uint64_t *subtree_sizes,
                    const uint64_t *parent_ids, const double *branches,
                    const double *rates, const double *T,
                    uint64_t burn_in, uint64_t batch, int do_reset,
                    double *J, double *D, double *acc_rate, double *seconds);
/* sharding-protocol test support (not reference functions): colour-class sweeps with
 * per-site reseeded mt19937, see oracle.c */
void orc_sampler_init(uint64_t n_nodes, const uint64_t *subtree_sizes,
                      const uint64_t *parent_ids, const double *branches,
                      const double *rates, const double *T);
uint64_t orc_colour_pass(int colour, uint64_t seed, uint64_t sweep, int64_t global0,
                         uint64_t n_global, uint64_t lo, uint64_t hi);
void orc_set_row(uint64_t site, uint64_t node, int init_state, uint64_t n, const double *jumps);
void orc_get_tri_cache(double *out);
void orc_set_tri_cache(uint64_t site, double v);
double orc_time_suff_stats(uint64_t reps, double *J, double *D);
uint64_t orc_end_cond_sample(double r0, double r1, int a, int b, double T,
                             uint64_t cap, double *jumps_out);

/* a14 */
double orc_log_likelihood(const double *J, const double *D, const double *rates);
void orc_gradient(const double *J, const double *D, const double *rates,
                  double *grad);
/* model helpers */
void orc_model_from_params(double T00, double T11, double b00, double b11,
                           int scale, double *rates, double *T, double *Q,
                           double *baseline);
void orc_model_rebuild(const double *rates_in, double *T, double *Q,
                       double *baseline);
double orc_rate_scaling_factor(const double *rates);
/* a15/a16/a17 */
double orc_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                          double param_tol, double *rates, double *T,
                          double *baseline);
double orc_estimate_rates_and_branches(uint64_t n_nodes, const double *J,
                                       const double *D, double param_tol,
                                       double *rates, double *branches,
                                       double *T, double *baseline);
void orc_set_one_change(double *rates, uint64_t n_nodes, double *branches);
void orc_scale_jump_times(uint64_t n_nodes, const double *new_branches);

#ifdef __cplusplus
}
#endif
#endif
