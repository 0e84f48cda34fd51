/* oracle/oracle.c -- TEST INFRASTRUCTURE ONLY.  See oracle/oracle.h.
 *
 * Plain-C (C99, libm only) restatement of the reference's MCMC /
 * sufficient-statistics hot path.  Floating-point expressions keep the
 * reference's association so that results are bit-identical with the
 * reference build on the same libm; each function cites the reference
 * file:line (relative to /root/reference/src/libepievo/) it follows.
 */
#define _POSIX_C_SOURCE 200809L
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define N_TRIPLES 8

/* ------------------------------------------------------------------------ */
/* growable double vector                                                    */
typedef struct { double *v; size_t n, cap; } dvec;

static void
dvec_push(dvec *a, double x) {
  if (a->n == a->cap) {
    a->cap = a->cap ? 2*a->cap : 4;
    a->v = (double *)realloc(a->v, a->cap*sizeof(double));
  }
  a->v[a->n++] = x;
}
static void dvec_clear(dvec *a) { a->n = 0; }
static void dvec_free(dvec *a) { free(a->v); a->v = 0; a->n = a->cap = 0; }
static void
dvec_assign(dvec *a, const double *src, size_t n) {
  dvec_clear(a);
  for (size_t i = 0; i < n; ++i) dvec_push(a, src[i]);
}

/* struct Path, Path.hpp:33-57 */
typedef struct { int init_state; double tot_time; dvec jumps; } path_t;

static int
end_state(const path_t *p) {                      /* Path.hpp:53-55 */
  return (p->jumps.n % 2 == 0) ? p->init_state : !p->init_state;
}

/* epievo_utils.hpp:85-115 */
static size_t triple2idx(int i, int j, int k) { return (size_t)(i*4 + j*2 + k); }
static size_t flip_left(size_t x) { return x ^ 4ul; }
static size_t flip_mid(size_t x) { return x ^ 2ul; }
static size_t flip_right(size_t x) { return x ^ 1ul; }

/* ------------------------------------------------------------------------ */
/* a3: ContinuousTimeMarkovModel.cpp:142-161                                 */
void
orc_trans_prob_mat(double r0, double r1, double t, double *P) {
  const double h = 1.0/exp(t*(r0 + r1));
  const double denominator = r0 + r1;
  P[0] = (r0*h + r1)/denominator;
  P[1] = 1.0 - P[0];
  P[3] = (r0 + r1*h)/denominator;
  P[2] = 1.0 - P[3];
}

/* a3: ContinuousTimeMarkovModel.cpp:115-125 */
double
orc_trans_prob(double r0, double r1, double t, int a, int b) {
  const double h = exp(-t*(r0 + r1));
  const double denominator = r0 + r1;
  const double prob = (a ? r0 + r1*h : r0*h + r1)/denominator;
  return (b == a) ? prob : 1.0 - prob;
}

/* ContinuousTimeMarkovModel.cpp:167-226 (known answers for a6) */
void
orc_expectations(double r0, double r1, double T,
                 double *J0, double *J1, double *D0, double *D1) {
  {
    const double s = r0 + r1, p = r0*r1, d = r1 - r0, e = exp(-s*T);
    const double C1 = d*(1 - e)/s;
    J0[0] = p*(T*(r1 - r0*e) - C1)/(s*(r1 + r0*e));
    J1[0] = J0[0];
    J0[3] = p*(T*(r0 - r1*e) + C1)/(s*(r0 + r1*e));
    J1[3] = J0[3];
    const double C2 = p*T*(1 + e)/(s*(1 - e));
    const double C3 = (r0*r0 + r1*r1)/(s*s);
    const double C4 = (2*p)/(s*s);
    J0[1] = C2 + C3;
    J1[1] = C2 - C4;
    J0[2] = J1[1];
    J1[2] = J0[1];
  }
  {
    const double r00 = r0*r0, r11 = r1*r1, s = r0 + r1, p = r0*r1;
    const double e = exp(-s*T);
    const double C1 = 2*p*(1 - e)/s;
    D0[0] = ((r11 + r00*e)*T + C1)/(s*(r1 + r0*e));
    D1[0] = T - D0[0];
    D1[3] = ((r00 + r11*e)*T + C1)/(s*(r0 + r1*e));
    D0[3] = T - D1[3];
    const double C2 = (p - r00)*(1 - e)/s;
    D1[1] = ((r00 - p*e)*T + C2)/(s*(r0 - r0*e));
    D0[1] = T - D1[1];
    const double C3 = (p - r11)*(1 - e)/s;
    D0[2] = ((r11 - p*e)*T + C3)/(s*(r1 - r1*e));
    D1[2] = T - D0[2];
  }
}

/* ------------------------------------------------------------------------ */
/* a2: Segment.hpp:32-42, Segment.cpp:35-79                                  */
typedef struct { double rate0, rate1; size_t trip0, trip1; double len; } seg_t;
typedef struct { seg_t *v; size_t n, cap; } segvec;

static void
seg_push(segvec *s, const double *rates, size_t t0, size_t t1, double len) {
  if (s->n == s->cap) {
    s->cap = s->cap ? 2*s->cap : 4;
    s->v = (seg_t *)realloc(s->v, s->cap*sizeof(seg_t));
  }
  seg_t x = { rates[t0], rates[t1], t0, t1, len };
  s->v[s->n++] = x;
}

static void
collect_segment_info(const double *rates, const path_t *l, const path_t *r,
                     segvec *seg) {
  seg->n = 0;
  size_t trip0 = triple2idx(l->init_state, 0, r->init_state);
  size_t trip1 = triple2idx(l->init_state, 1, r->init_state);
  double prev_time = 0.0;
  size_t i = 0, j = 0;
  const size_t i_lim = l->jumps.n, j_lim = r->jumps.n;
  const double *lj = l->jumps.v, *rj = r->jumps.v;
  while (i != i_lim && j != j_lim)
    if (lj[i] < rj[j]) {                       /* strict: ties go right first */
      seg_push(seg, rates, trip0, trip1, lj[i] - prev_time);
      trip0 = flip_left(trip0); trip1 = flip_left(trip1);
      prev_time = lj[i++];
    }
    else {
      seg_push(seg, rates, trip0, trip1, rj[j] - prev_time);
      trip0 = flip_right(trip0); trip1 = flip_right(trip1);
      prev_time = rj[j++];
    }
  for (; i != i_lim; ++i) {
    seg_push(seg, rates, trip0, trip1, lj[i] - prev_time);
    trip0 = flip_left(trip0); trip1 = flip_left(trip1);
    prev_time = lj[i];
  }
  for (; j != j_lim; ++j) {
    seg_push(seg, rates, trip0, trip1, rj[j] - prev_time);
    trip0 = flip_right(trip0); trip1 = flip_right(trip1);
    prev_time = rj[j];
  }
  seg_push(seg, rates, trip0, trip1, l->tot_time - prev_time);
}

static path_t
view_path(int init, double tot, uint64_t n, const double *j) {
  path_t p;
  p.init_state = init != 0; p.tot_time = tot;
  p.jumps.v = (double *)j; p.jumps.n = n; p.jumps.cap = n;
  return p;
}

uint64_t
orc_segment_info(const double *rates, int l_init, uint64_t nl, const double *lj,
                 int r_init, uint64_t nr, const double *rj, double tot_time,
                 double *rate0, double *rate1, uint64_t *trip0, uint64_t *trip1,
                 double *len) {
  const path_t l = view_path(l_init, tot_time, nl, lj);
  const path_t r = view_path(r_init, tot_time, nr, rj);
  segvec s = {0, 0, 0};
  collect_segment_info(rates, &l, &r, &s);
  for (size_t i = 0; i < s.n; ++i) {
    rate0[i] = s.v[i].rate0; rate1[i] = s.v[i].rate1;
    trip0[i] = s.v[i].trip0; trip1[i] = s.v[i].trip1; len[i] = s.v[i].len;
  }
  const uint64_t n = s.n;
  free(s.v);
  return n;
}

/* ------------------------------------------------------------------------ */
/* a12: Path.cpp:206-301                                                     */
static void
add_suff_stat_single(size_t change_pos, const path_t *a, size_t i,
                     size_t triplet, double prev_time, double *J, double *D) {
  while (i < a->jumps.n) {                                   /* :209-218 */
    D[triplet] += a->jumps.v[i] - prev_time;
    if (change_pos == 1) J[triplet] += 1.0;
    triplet = (change_pos == 1) ? flip_mid(triplet) :
      (change_pos == 0) ? flip_left(triplet) : flip_right(triplet);
    prev_time = a->jumps.v[i];
    ++i;
  }
  D[triplet] += a->tot_time - prev_time;                     /* :221 */
}

static void
add_suff_stat_pair(size_t fixed_pos, const path_t *a, size_t i,
                   const path_t *b, size_t j, size_t triplet, double prev_time,
                   double *J, double *D) {
  while (i < a->jumps.n && j < b->jumps.n)                   /* :244-261 */
    if (a->jumps.v[i] < b->jumps.v[j]) {
      D[triplet] += a->jumps.v[i] - prev_time;
      if (fixed_pos == 0) J[triplet] += 1.0;
      triplet = fixed_pos == 0 ? flip_mid(triplet) : flip_left(triplet);
      prev_time = a->jumps.v[i];
      ++i;
    }
    else {
      D[triplet] += b->jumps.v[j] - prev_time;
      if (fixed_pos == 2) J[triplet] += 1.0;
      triplet = fixed_pos == 2 ? flip_mid(triplet) : flip_right(triplet);
      prev_time = b->jumps.v[j];
      ++j;
    }
  if (i < a->jumps.n)                                        /* :262-265 */
    add_suff_stat_single(fixed_pos == 0 ? 1 : 0, a, i, triplet, prev_time, J, D);
  else
    add_suff_stat_single(fixed_pos == 2 ? 1 : 2, b, j, triplet, prev_time, J, D);
}

static void
add_sufficient_statistics(const path_t *left, const path_t *mid,
                          const path_t *right, double *J, double *D) {
  size_t triplet = triple2idx(left->init_state, mid->init_state,
                              right->init_state);
  double prev_time = 0.0;
  size_t i = 0, j = 0, k = 0;
  const double *lj = left->jumps.v, *mj = mid->jumps.v, *rj = right->jumps.v;
  while (i < left->jumps.n && j < mid->jumps.n && k < right->jumps.n) {
    const double mn = mj[j] < rj[k] ? mj[j] : rj[k];   /* std::min(mid, right) */
    if (lj[i] < mn) {                                         /* LEFT :275 */
      D[triplet] += lj[i] - prev_time;
      prev_time = lj[i];
      triplet = flip_left(triplet);
      ++i;
    }
    else if (mj[j] < rj[k]) {                                 /* MID :281 */
      D[triplet] += mj[j] - prev_time;
      J[triplet] += 1.0;
      prev_time = mj[j];
      triplet = flip_mid(triplet);
      ++j;
    }
    else {                                                    /* RIGHT :288 */
      D[triplet] += rj[k] - prev_time;
      prev_time = rj[k];
      triplet = flip_right(triplet);
      ++k;
    }
  }
  if (i == left->jumps.n)                                     /* :295-300 */
    add_suff_stat_pair(0, mid, j, right, k, triplet, prev_time, J, D);
  else if (j == mid->jumps.n)
    add_suff_stat_pair(1, left, i, right, k, triplet, prev_time, J, D);
  else
    add_suff_stat_pair(2, left, i, mid, j, triplet, prev_time, J, D);
}

void
orc_add_suff_stats(int l_init, uint64_t nl, const double *lj,
                   int m_init, uint64_t nm, const double *mj,
                   int r_init, uint64_t nr, const double *rj,
                   double tot_time, double *J, double *D) {
  const path_t l = view_path(l_init, tot_time, nl, lj);
  const path_t m = view_path(m_init, tot_time, nm, mj);
  const path_t r = view_path(r_init, tot_time, nr, rj);
  add_sufficient_statistics(&l, &m, &r, J, D);
}

/* ------------------------------------------------------------------------ */
/* global path container: paths[site][node], as the reference's callers hold it
 * (epievo_est_params_histories.cpp:180-192)                                  */
static void sampler_free(void);
static path_t *g_paths = 0;
static size_t g_n_sites = 0, g_n_nodes = 0;
#define PATH(s, b) (g_paths[(s)*g_n_nodes + (b)])

void
orc_free_paths(void) {
  for (size_t i = 0; i < g_n_sites*g_n_nodes; ++i) dvec_free(&g_paths[i].jumps);
  free(g_paths);
  g_paths = 0; g_n_sites = g_n_nodes = 0;
}

void
orc_load_paths(uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
               const uint64_t *row_off, const double *jumps,
               const double *tot_time) {
  sampler_free();
  orc_free_paths();
  g_n_sites = n_sites; g_n_nodes = n_nodes;
  g_paths = (path_t *)calloc(n_sites*n_nodes, sizeof(path_t));
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t s = 0; s < n_sites; ++s) {
      const size_t row = b*n_sites + s;
      path_t *p = &PATH(s, b);
      p->init_state = init_state[row] != 0;
      p->tot_time = tot_time[b];
      dvec_assign(&p->jumps, jumps + row_off[row], row_off[row+1] - row_off[row]);
    }
}

/* overwrite one row in place (halo refresh in the sharding-protocol tests) */
void
orc_set_row(uint64_t site, uint64_t node, int init_state, uint64_t n, const double *jumps) {
  path_t *p = &PATH(site, node);
  p->init_state = init_state != 0;
  dvec_assign(&p->jumps, jumps, n);
}

uint64_t
orc_total_jumps(void) {
  uint64_t n = 0;
  for (size_t i = 0; i < g_n_sites*g_n_nodes; ++i) n += g_paths[i].jumps.n;
  return n;
}

void
orc_store_paths(uint8_t *init_state, uint64_t *row_off, double *jumps,
                double *tot_time) {
  uint64_t off = 0;
  for (size_t b = 0; b < g_n_nodes; ++b) {
    tot_time[b] = (b && g_n_sites) ? PATH(0, b).tot_time : 0.0;
    for (size_t s = 0; s < g_n_sites; ++s) {
      const size_t row = b*g_n_sites + s;
      const path_t *p = &PATH(s, b);
      init_state[row] = (uint8_t)p->init_state;
      row_off[row] = off;
      for (size_t k = 0; k < p->jumps.n; ++k) jumps[off++] = p->jumps.v[k];
    }
  }
  row_off[g_n_nodes*g_n_sites] = off;
}

/* a13: ParamEstimation.cpp:92-114 (site-major summation order) */
void
orc_get_suff_stats(double *J, double *D) {
  for (size_t i = 0; i < g_n_nodes*N_TRIPLES; ++i) J[i] = D[i] = 0.0;
  for (size_t s = 1; s + 1 < g_n_sites; ++s)
    for (size_t b = 1; b < g_n_nodes; ++b)
      add_sufficient_statistics(&PATH(s-1, b), &PATH(s, b), &PATH(s+1, b),
                                J + b*N_TRIPLES, D + b*N_TRIPLES);
}

double
orc_time_suff_stats(uint64_t reps, double *J, double *D) {
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (uint64_t r = 0; r < reps; ++r) orc_get_suff_stats(J, D);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9*(double)(t1.tv_nsec - t0.tv_nsec);
}

/* ------------------------------------------------------------------------ */
/* a8: SingleSiteSampler.cpp:263-269, :342-352, :374-391.  NB the root term is
 * the raw probability T(l,m)*T(m,r), not its log -- a reference quirk that
 * cancels in every acceptance ratio; reproduced for parity.                 */
static double
path_log_likelihood(const double *rates, const double *log_rates, const double *T,
                    const path_t *l, const path_t *m, const path_t *r,
                    size_t n_nodes) {
  double J[N_TRIPLES], D[N_TRIPLES];
  for (size_t i = 0; i < N_TRIPLES; ++i) J[i] = D[i] = 0.0;
  double llh = T[l[1].init_state*2 + m[1].init_state]*
    T[m[1].init_state*2 + r[1].init_state];
  for (size_t i = 1; i < n_nodes; ++i)
    add_sufficient_statistics(&l[i], &m[i], &r[i], J, D);
  double x = 0.0;
  for (size_t i = 0; i < N_TRIPLES; ++i)
    x += J[i]*log_rates[i] - D[i]*rates[i];
  llh += x;
  return llh;
}

/* a10: SingleSiteSampler.cpp:464-474 */
void
orc_tri_llh(const double *rates, const double *T, double *tri_llh) {
  double log_rates[N_TRIPLES];
  for (size_t i = 0; i < N_TRIPLES; ++i) log_rates[i] = log(rates[i]);
  for (size_t s = 0; s < g_n_sites; ++s) tri_llh[s] = 0.0;
  for (size_t s = 1; s + 1 < g_n_sites; ++s)
    tri_llh[s] = path_log_likelihood(rates, log_rates, T, &PATH(s-1, 0),
                                     &PATH(s, 0), &PATH(s+1, 0), g_n_nodes);
}

/* ------------------------------------------------------------------------ */
/* a4: FelsHelper SingleSiteSampler.cpp:65-71; pruning :80-157               */
typedef struct { dvec p; double q[2]; } fels_t;   /* p holds pairs */

static int
is_leaf(const uint64_t *subtree_sizes, size_t node) {
  return subtree_sizes[node] == 1;                 /* PhyloTreePreorder.hpp */
}

static void
process_interval(const seg_t *si, const double *q, double *p) {  /* :80-88 */
  double P[4];
  orc_trans_prob_mat(si->rate0, si->rate1, si->len, P);
  const double p0 = P[0]*q[0] + P[1]*q[1];
  const double p1 = P[2]*q[0] + P[3]*q[1];
  p[0] = p0; p[1] = p1;
}

static void
pruning(size_t n_nodes, const uint64_t *subtree_sizes, size_t site,
        const segvec *seg, fels_t *fh) {
  for (size_t i = n_nodes; i > 0; --i) {                    /* :153-156 */
    const size_t node = i - 1;
    double q[2] = {1.0, 1.0};
    if (is_leaf(subtree_sizes, node)) {                     /* :122-126 */
      const int leaf_state = end_state(&PATH(site, node));
      q[0] = (leaf_state == 0) ? 1.0 : 0.0;
      q[1] = (leaf_state == 1) ? 1.0 : 0.0;
    }
    else {                                                  /* ChildSet :128-131 */
      for (size_t ch = 1; ch < subtree_sizes[node];
           ch += subtree_sizes[node + ch]) {
        q[0] *= fh[node + ch].p.v[0];
        q[1] *= fh[node + ch].p.v[1];
      }
    }
    fh[node].q[0] = q[0]; fh[node].q[1] = q[1];
    if (node != 0) {                                        /* :91-106 */
      const size_t n_int = seg[node].n;
      dvec_clear(&fh[node].p);
      for (size_t k = 0; k < 2*n_int; ++k) dvec_push(&fh[node].p, 0.0);
      double *p = fh[node].p.v;
      process_interval(&seg[node].v[n_int-1], fh[node].q, p + 2*(n_int-1));
      for (size_t k = n_int - 1; k > 0; --k)
        process_interval(&seg[node].v[k-1], p + 2*k, p + 2*(k-1));
    }
  }
}

/* :167-176 */
static double
root_post_prob0(int left_st, int right_st, const double *T, const double *q) {
  const double p0 = (T[left_st*2 + 0]*T[0*2 + right_st])*q[0];
  const double p1 = (T[left_st*2 + 1]*T[1*2 + right_st])*q[1];
  return p0/(p0 + p1);
}

/* a7: SingleSiteSampler.cpp:272-312 */
static double
proposal_prob_branch(const segvec *seg, const fels_t *fh, const path_t *path) {
  const size_t n_int = seg->n;
  size_t start_state = (size_t)path->init_state, end_st = (size_t)path->init_state;
  double end_time = 0.0;
  size_t start_jump = 0, end_jump = 0;
  double log_prob = 0.0;
  for (size_t i = 0; i < n_int; ++i) {
    end_time += seg->v[i].len;
    while (end_jump < path->jumps.n && path->jumps.v[end_jump] < end_time)
      ++end_jump;
    if ((end_jump - start_jump) % 2 == 1) end_st = 1 - end_st;
    const double r0 = seg->v[i].rate0, r1 = seg->v[i].rate1;
    const double PT0 = orc_trans_prob(r0, r1, seg->v[i].len, (int)start_state, 0);
    log_prob -= log(orc_trans_prob(r0, r1, seg->v[i].len, (int)start_state,
                                   (int)end_st));
    const double p0 = PT0/fh->p.v[2*i + start_state]*
      ((i == n_int - 1) ? fh->q[0] : fh->p.v[2*(i+1)]);
    log_prob += (end_st == 0) ? log(p0) : log(1.0 - p0);
    start_jump = end_jump;
    start_state = end_st;
  }
  return log_prob;
}

/* :316-339 with sample_root == false (it always is, :441) */
static double
proposal_prob(size_t n_nodes, const fels_t *fh, const segvec *seg,
              const path_t *the_path) {
  double log_prob = 0.0;
  for (size_t node = 1; node < n_nodes; ++node)
    log_prob += proposal_prob_branch(&seg[node], &fh[node], &the_path[node]);
  return log_prob;
}

uint64_t
orc_site_debug(uint64_t n_nodes, const uint64_t *subtree_sizes,
               const uint64_t *parent_ids, const double *branches,
               const double *rates, const double *T, uint64_t site,
               uint64_t cap, uint64_t *seg_off, double *rate0, double *rate1,
               uint64_t *trip0, uint64_t *trip1, double *len,
               double *p, double *q, double *root_p0, double *proposal_logp) {
  (void)parent_ids; (void)branches;
  segvec *seg = (segvec *)calloc(n_nodes, sizeof(segvec));
  fels_t *fh = (fels_t *)calloc(n_nodes, sizeof(fels_t));
  for (size_t b = 1; b < n_nodes; ++b)
    collect_segment_info(rates, &PATH(site-1, b), &PATH(site+1, b), &seg[b]);
  pruning(n_nodes, subtree_sizes, site, seg, fh);
  uint64_t k = 0, ret = 0;
  seg_off[0] = 0;
  for (size_t b = 0; b < n_nodes && ret != UINT64_MAX; ++b) {
    for (size_t i = 0; i < seg[b].n; ++i, ++k) {
      if (k >= cap) { ret = UINT64_MAX; break; }
      rate0[k] = seg[b].v[i].rate0; rate1[k] = seg[b].v[i].rate1;
      trip0[k] = seg[b].v[i].trip0; trip1[k] = seg[b].v[i].trip1;
      len[k] = seg[b].v[i].len;
      p[2*k] = fh[b].p.v[2*i]; p[2*k+1] = fh[b].p.v[2*i+1];
    }
    seg_off[b+1] = k;
    q[2*b] = fh[b].q[0]; q[2*b+1] = fh[b].q[1];
  }
  if (ret != UINT64_MAX) {
    const int ls = PATH(site-1, 1).init_state, rs = PATH(site+1, 1).init_state;
    *root_p0 = root_post_prob0(ls, rs, T, fh[0].q);
    *proposal_logp = proposal_prob(n_nodes, fh, seg, &PATH(site, 0));
    ret = k;
  }
  for (size_t b = 0; b < n_nodes; ++b) { free(seg[b].v); dvec_free(&fh[b].p); }
  free(seg); free(fh);
  return ret;
}

/* ------------------------------------------------------------------------ */
/* std::mt19937 and the libstdc++ distributions the reference draws from
 * (published MT19937 algorithm; <bits/random.tcc>: generate_canonical<double,
 * 53> takes two 32-bit draws, low word first; exponential_distribution is
 * -log(1-u)/lambda; uniform_real_distribution(0,1) is u).                   */
static uint32_t mt[624];
static int mti = 625;

void
orc_seed(uint64_t seed) {
  mt[0] = (uint32_t)seed;
  for (int i = 1; i < 624; ++i)
    mt[i] = 1812433253u*(mt[i-1] ^ (mt[i-1] >> 30)) + (uint32_t)i;
  mti = 624;
}

static uint32_t
mt_next(void) {
  if (mti >= 624) {
    if (mti == 625) orc_seed(5489u);
    for (int k = 0; k < 624; ++k) {
      const uint32_t y = (mt[k] & 0x80000000u) | (mt[(k+1) % 624] & 0x7fffffffu);
      mt[k] = mt[(k+397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    mti = 0;
  }
  uint32_t y = mt[mti++];
  y ^= (y >> 11);
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= (y >> 18);
  return y;
}

static double
canonical(void) {
  double sum = 0.0, tmp = 1.0;
  for (int k = 0; k < 2; ++k) {
    sum += (double)mt_next()*tmp;
    tmp *= 4294967296.0;
  }
  double ret = sum/tmp;
  if (ret >= 1.0) ret = nextafter(1.0, 0.0);
  return ret;
}

static double rexp(double lambda) { return -log(1.0 - canonical())/lambda; }

/* a6: EndCondSampling.cpp:466-476 */
static size_t
forward_sampling(const double *lambda, size_t a, double T, double start_time,
                 dvec *jump_times) {
  double tau = start_time;
  while ((tau += rexp(lambda[a])) < T) {
    a = 1 - a;
    dvec_push(jump_times, tau);
  }
  return a;
}

/* a6: EndCondSampling.cpp:479-509 (give-up cap 1e10, :52) */
static int
end_cond_sample_forward_rejection(double rate0, double rate1, size_t start_state,
                                  size_t end_st, double T, dvec *jump_times,
                                  double start_time, dvec *proposal) {
  const double lambda[2] = { rate0, rate1 };
  const uint64_t max_sample_count = 10000000000ull;
  uint64_t sample_count = 0;
  dvec_clear(proposal);
  while (forward_sampling(lambda, start_state, T, 0.0, proposal) != end_st &&
         sample_count < max_sample_count) {
    ++sample_count;
    dvec_clear(proposal);
  }
  if (sample_count < max_sample_count)
    for (size_t k = 0; k < proposal->n; ++k)
      dvec_push(jump_times, proposal->v[k] + start_time);
  return sample_count < max_sample_count;
}

uint64_t
orc_end_cond_sample(double r0, double r1, int a, int b, double T,
                    uint64_t cap, double *jumps_out) {
  dvec j = {0, 0, 0}, scratch = {0, 0, 0};
  end_cond_sample_forward_rejection(r0, r1, (size_t)a, (size_t)b, T, &j, 0.0,
                                    &scratch);
  const uint64_t n = j.n;
  for (size_t k = 0; k < n && k < cap; ++k) jumps_out[k] = j.v[k];
  dvec_free(&j); dvec_free(&scratch);
  return n;
}

/* ------------------------------------------------------------------------ */
/* the sampler: SingleSiteSampler.cpp:180-255 (downward), :396-433 (ratio),
 * :482-536 (MH site), :538-598 (sweep, run_mcmc)                            */
typedef struct {
  size_t n_nodes;
  const uint64_t *subtree_sizes, *parent_ids;
  const double *branches, *rates, *T;
  double log_rates[N_TRIPLES];
  double *tri_llh;
  path_t *proposed;      /* [n_nodes] */
  segvec *seg;           /* [n_nodes] */
  fels_t *fh;            /* [n_nodes] */
  dvec scratch;
} sampler_t;

static void
downward_sampling_branch(sampler_t *S, size_t node, size_t start_state,
                         double *log_prob) {
  const segvec *seg = &S->seg[node];
  const fels_t *fh = &S->fh[node];
  path_t *sp = &S->proposed[node];
  sp->init_state = (int)start_state;                         /* :190-192 */
  sp->tot_time = S->branches[node];
  dvec_clear(&sp->jumps);
  size_t prev_state = start_state;
  double time_passed = 0.0;
  const size_t n_int = seg->n;
  for (size_t i = 0; i < n_int; ++i) {                       /* :197-220 */
    const double r0 = seg->v[i].rate0, r1 = seg->v[i].rate1;
    const double PT0 = orc_trans_prob(r0, r1, seg->v[i].len, (int)prev_state, 0);
    const double p0 = PT0*((i == n_int - 1) ? fh->q[0] : fh->p.v[2*(i+1)])/
      fh->p.v[2*i + prev_state];
    const size_t sampled_state = (canonical() > p0);
    *log_prob += (sampled_state == 0) ? log(p0) : log(1.0 - p0);
    /* exp_distr[trip] carries rates[trip] (:461-462), which is rate0/rate1 */
    end_cond_sample_forward_rejection(S->rates[seg->v[i].trip0],
                                      S->rates[seg->v[i].trip1],
                                      prev_state, sampled_state, seg->v[i].len,
                                      &sp->jumps, time_passed, &S->scratch);
    *log_prob -= log(orc_trans_prob(r0, r1, seg->v[i].len, (int)prev_state,
                                    (int)sampled_state));
    time_passed += seg->v[i].len;
    prev_state = sampled_state;
  }
}

static int
metropolis_hastings_site(sampler_t *S, size_t site) {
  const size_t n_nodes = S->n_nodes;
  for (size_t b = 1; b < n_nodes; ++b)                       /* :489-494 */
    collect_segment_info(S->rates, &PATH(site-1, b), &PATH(site+1, b),
                         &S->seg[b]);
  pruning(n_nodes, S->subtree_sizes, site, S->seg, S->fh);  /* :498 */

  double update_proposal = 0.0;                              /* :500-507 */
  S->proposed[0].init_state = PATH(site, 1).init_state;     /* :246 */
  for (size_t node = 1; node < n_nodes; ++node) {           /* :249-254 */
    const size_t start_state =
      (size_t)end_state(&S->proposed[S->parent_ids[node]]);
    downward_sampling_branch(S, node, start_state, &update_proposal);
  }

  double llh_l = S->tri_llh[site-1];                         /* :510-512 */
  double llh_m = S->tri_llh[site];
  double llh_r = S->tri_llh[site+1];

  /* log_accept_rate :396-433 */
  const double orig_proposal = proposal_prob(n_nodes, S->fh, S->seg,
                                             &PATH(site, 0));
  double llr = orig_proposal - update_proposal;
  const double llh_l_orig = llh_l, llh_r_orig = llh_r;
  if (site > 1)
    llh_l = path_log_likelihood(S->rates, S->log_rates, S->T, &PATH(site-2, 0),
                                &PATH(site-1, 0), S->proposed, n_nodes);
  llh_m = path_log_likelihood(S->rates, S->log_rates, S->T, &PATH(site-1, 0),
                              S->proposed, &PATH(site+1, 0), n_nodes);
  if (site < g_n_sites - 2)
    llh_r = path_log_likelihood(S->rates, S->log_rates, S->T, S->proposed,
                                &PATH(site+1, 0), &PATH(site+2, 0), n_nodes);
  llr += (llh_l + llh_r - llh_l_orig - llh_r_orig);

  const double u = canonical();                              /* :520-525 */
  int accepted = 0;
  if (llr >= 0 || u < exp(llr)) accepted = 1;
  if (accepted) {                                            /* :528-533 */
    for (size_t b = 0; b < n_nodes; ++b) {
      path_t tmp = PATH(site, b);
      PATH(site, b) = S->proposed[b];
      S->proposed[b] = tmp;
    }
    S->tri_llh[site-1] = llh_l;
    S->tri_llh[site] = llh_m;
    S->tri_llh[site+1] = llh_r;
  }
  return accepted;
}

static size_t
single_iteration(sampler_t *S) {                             /* :538-548 */
  size_t n_accepted = 0;
  for (size_t site = 1; site + 1 < g_n_sites; ++site)
    n_accepted += (size_t)metropolis_hastings_site(S, site);
  return n_accepted;
}

static sampler_t g_S;
static int g_S_live = 0;

static void
sampler_free(void) {
  if (!g_S_live) return;
  for (size_t b = 0; b < g_S.n_nodes; ++b) {
    dvec_free(&g_S.proposed[b].jumps);
    free(g_S.seg[b].v);
    dvec_free(&g_S.fh[b].p);
  }
  free(g_S.proposed); free(g_S.seg); free(g_S.fh); free(g_S.tri_llh);
  dvec_free(&g_S.scratch);
  g_S_live = 0;
}

static void
sampler_init(uint64_t n_nodes, const double *rates, const double *T) {  /* reset :449-475 */
  sampler_free();
  memset(&g_S, 0, sizeof g_S);
  g_S.n_nodes = n_nodes;
  g_S.proposed = (path_t *)calloc(n_nodes, sizeof(path_t));
  g_S.seg = (segvec *)calloc(n_nodes, sizeof(segvec));
  g_S.fh = (fels_t *)calloc(n_nodes, sizeof(fels_t));
  g_S.tri_llh = (double *)calloc(g_n_sites, sizeof(double));
  for (size_t i = 0; i < N_TRIPLES; ++i) g_S.log_rates[i] = log(rates[i]);
  orc_tri_llh(rates, T, g_S.tri_llh);
  g_S_live = 1;
}

/* ---- colour-class sweeps with per-site reseeding (TEST INFRASTRUCTURE for the
 * sharding protocol; not a reference function).  Each update of global site g in sweep
 * k draws from mt19937 reseeded with a hash of (seed, k, g), so the result of a colour
 * pass does not depend on how the sites are partitioned into shards -- the property the
 * GPU path gets from its (seed, sweep, site)-keyed Philox streams. */
void
orc_sampler_init(uint64_t n_nodes, const uint64_t *subtree_sizes,
                 const uint64_t *parent_ids, const double *branches,
                 const double *rates, const double *T) {
  sampler_init(n_nodes, rates, T);
  g_S.subtree_sizes = subtree_sizes; g_S.parent_ids = parent_ids;
  g_S.branches = branches; g_S.rates = rates; g_S.T = T;
}

uint64_t
orc_colour_pass(int colour, uint64_t seed, uint64_t sweep, int64_t global0,
                uint64_t n_global, uint64_t lo, uint64_t hi) {
  uint64_t n_accepted = 0;
  for (uint64_t s = lo; s < hi; ++s) {
    const int64_t g = global0 + (int64_t)s;
    if ((int)(s % 3) != colour || g < 1 || g + 1 >= (int64_t)n_global) continue;
    uint64_t h = seed*0x9E3779B97F4A7C15ull + sweep;
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h += (uint64_t)g;
    h ^= h >> 32; h *= 0x94D049BB133111EBull; h ^= h >> 29;
    orc_seed(h & 0xffffffffu);
    n_accepted += (uint64_t)metropolis_hastings_site(&g_S, s);
  }
  return n_accepted;
}

void orc_get_tri_cache(double *out) { memcpy(out, g_S.tri_llh, g_n_sites*sizeof(double)); }
void orc_set_tri_cache(uint64_t site, double v) { g_S.tri_llh[site] = v; }

double
orc_run_mcmc(uint64_t n_nodes, const uint64_t *subtree_sizes,
             const uint64_t *parent_ids, const double *branches,
             const double *rates, const double *T,
             uint64_t burn_in, uint64_t batch, int do_reset,
             double *J, double *D, double *acc_rate, double *seconds) {
  if (!g_S_live || do_reset) sampler_init(n_nodes, rates, T);
  g_S.subtree_sizes = subtree_sizes; g_S.parent_ids = parent_ids;
  g_S.branches = branches; g_S.rates = rates; g_S.T = T;

  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (uint64_t it = 0; it < burn_in; ++it) single_iteration(&g_S); /* :560 */
  for (size_t i = 0; i < n_nodes*N_TRIPLES; ++i) J[i] = D[i] = 0.0;
  double *J1 = (double *)calloc(n_nodes*N_TRIPLES, sizeof(double));
  double *D1 = (double *)calloc(n_nodes*N_TRIPLES, sizeof(double));
  size_t n_accepted = 0;
  for (uint64_t it = 0; it < batch; ++it) {                  /* :576-587 */
    n_accepted += single_iteration(&g_S);
    orc_get_suff_stats(J1, D1);
    for (size_t b = 1; b < n_nodes; ++b)
      for (size_t i = 0; i < N_TRIPLES; ++i) {
        J[b*N_TRIPLES + i] += J1[b*N_TRIPLES + i];
        D[b*N_TRIPLES + i] += D1[b*N_TRIPLES + i];
      }
  }
  for (size_t b = 1; b < n_nodes; ++b)                       /* :590-594 */
    for (size_t i = 0; i < N_TRIPLES; ++i) {
      J[b*N_TRIPLES + i] /= batch;
      D[b*N_TRIPLES + i] /= batch;
    }
  *acc_rate = (double)n_accepted/(batch*(g_n_sites - 2));    /* :597 */
  clock_gettime(CLOCK_MONOTONIC, &t1);
  free(J1); free(D1);
  const double sec =
    (double)(t1.tv_sec - t0.tv_sec) + 1e-9*(double)(t1.tv_nsec - t0.tv_nsec);
  if (seconds) *seconds = sec;
  return sec;
}

/* ------------------------------------------------------------------------ */
/* a14: ParamEstimation.cpp:131-143 */
double
orc_log_likelihood(const double *J, const double *D, const double *rates) {
  double ll = 0;
  for (size_t i = 0; i < N_TRIPLES; ++i)
    ll += J[i]*log(rates[i]) - D[i]*rates[i];
  return ll;
}

/* a14: ParamEstimation.cpp:147-194 (gradient w.r.t. log-rates, zero-initialised) */
void
orc_gradient(const double *J, const double *D, const double *rates, double *g) {
  for (size_t i = 0; i < N_TRIPLES; ++i) g[i] = 0.0;
  const double factor_111 = (J[7] - D[7]*rates[7]);
  g[0] += J[0] - D[0]*rates[0] + factor_111;
  g[2] += J[2] - D[2]*rates[2] - factor_111;
  g[1] += J[1] + J[4] - (D[1] + D[4])*rates[1] - 2*factor_111;
  g[4] = g[1];
  g[3] += J[3] + J[6] - (D[3] + D[6])*rates[3] + 2*factor_111;
  g[6] = g[3];
  g[5] += J[5] - D[5]*rates[5] + factor_111;
}

/* ------------------------------------------------------------------------ */
/* EpiEvoModel.cpp:86-96, :102-113, :116-137, :151-156                       */
static void
T_to_Q(const double *T, double *Q) {
  Q[0] = 1.0 - T[1];
  Q[1] = sqrt(T[1]*T[2]);
  Q[2] = Q[1];
  Q[3] = 1.0 - T[2];
}

static void
rates_to_Qprop(const double *r, double *Q) {
  Q[0] = Q[1] = Q[2] = Q[3] = 1.0;
  const double death_birth_ratio = r[2]/r[0];
  const double expand_contract_ratio = r[1]/r[3];
  Q[0] = Q[1]*sqrt(death_birth_ratio);
  Q[3] = Q[1]*sqrt(death_birth_ratio)*expand_contract_ratio;
}

static void
Q_to_T(const double *Q, double *T) {
  const double delta = sqrt(pow(Q[0] - Q[3], 2) + 4*Q[1]*Q[2]);
  const double diag_denom = Q[0] + Q[3] + delta;
  T[3] = 2*Q[3]/diag_denom;
  T[0] = 2*Q[0]/diag_denom;
  T[1] = 1.0 - T[0];
  T[2] = 1.0 - T[3];
}

/* EpiEvoModel.cpp:158-189 */
double
orc_rate_scaling_factor(const double *rates) {
  double Qp[4], T[4], pi[2];
  rates_to_Qprop(rates, Qp);
  Q_to_T(Qp, T);
  pi[1] = (1.0 - T[0])/(2.0 - T[0] - T[3]);
  pi[0] = 1.0 - pi[1];
  double mu = 0.0;
  for (size_t i = 0; i < 8; ++i) {
    const size_t l = (i >> 2) & 1, m = (i >> 1) & 1, r = i & 1;
    mu += pi[l]*T[l*2 + m]*T[m*2 + r]*rates[i];
  }
  return mu;
}

/* EpiEvoModel.cpp:420-449 */
void
orc_model_rebuild(const double *rates, double *T, double *Q, double *bl) {
  double Qp[4];
  rates_to_Qprop(rates, Qp);
  Q_to_T(Qp, T);
  T_to_Q(T, Q);
  const double log_Q01 = log(Q[1]), log_Q10 = log(Q[2]), log_Q11 = log(Q[3]);
  bl[0] = log(rates[0]) - (log_Q01 + log_Q10);
  bl[1] = log(rates[1]) - (log_Q01 + log_Q11);
  bl[2] = log(rates[4]) - (log_Q11 + log_Q10);
  bl[3] = log(rates[7]) - (log_Q10 + log_Q01);
  const double c = bl[1];
  bl[0] -= c; bl[1] -= c; bl[2] -= c; bl[3] -= c;
}

/* read_model EpiEvoModel.cpp:319-340, initialize :379-392,
 * compute_triplet_rates :398-417, scale_triplet_rates :372-377 */
void
orc_model_from_params(double T00, double T11, double b00, double b11, int scale,
                      double *rates, double *T, double *Q, double *bl) {
  T[0] = T00; T[3] = T11;
  T[2] = 1.0 - T[3];
  T[1] = 1.0 - T[0];
  bl[0] = b00; bl[1] = 0.0; bl[2] = 0.0; bl[3] = b11;
  T_to_Q(T, Q);
  double e = exp(bl[0]);
  rates[0] = Q[1]*Q[2]*e;
  rates[2] = Q[0]*Q[0]*e;
  e = exp(bl[1]);
  rates[1] = Q[1]*Q[3]*e;
  rates[3] = Q[0]*Q[1]*e;
  e = exp(bl[2]);
  rates[4] = Q[3]*Q[2]*e;
  rates[6] = Q[2]*Q[0]*e;
  e = exp(bl[3]);
  rates[5] = Q[3]*Q[3]*e;
  rates[7] = Q[2]*Q[1]*e;
  if (scale) {
    const double mu = orc_rate_scaling_factor(rates);
    for (size_t i = 0; i < 8; ++i) rates[i] /= mu;
  }
}

/* ------------------------------------------------------------------------ */
/* a15: ParamEstimation.cpp:200-218 */
static void
candidate_rates(double step, const double *g, const double *rates, double *u) {
  for (size_t i = 0; i < 7; ++i) u[i] = exp(log(rates[i]) + g[i]*step);
  u[7] = exp(log(u[0]) + log(u[5]) + 2*log(u[3]) - log(u[2]) - 2*log(u[1]));
}

/* :243-249, :256-276 */
static int
gradient_ascent(double tol, const double *J, const double *D, double llh,
                const double *rates, double *updated_llh, double *updated) {
  double g[8];
  orc_gradient(J, D, rates, g);
  double l1 = 0.0;
  for (size_t i = 0; i < 8; ++i) l1 += fabs(g[i]);
  double step = 1.0/l1;
  *updated_llh = -DBL_MAX;
  while (*updated_llh < llh && step > tol) {
    candidate_rates(step, g, rates, updated);
    *updated_llh = orc_log_likelihood(J, D, updated);
    step *= 0.5;
  }
  return *updated_llh > llh;
}

/* :279-296 (8-vector), :299-315 (collapse over branches) */
static double
estimate_rates8(double tol, const double *J, const double *D,
                const double *input, double *rates) {
  double llh = orc_log_likelihood(J, D, input);
  double a[8], b[8], *cur = a, *tmp = b;
  memcpy(cur, input, sizeof a);
  memcpy(tmp, input, sizeof b);
  double tmp_llh = llh;
  while (gradient_ascent(tol, J, D, llh, cur, &tmp_llh, tmp)) {
    llh = tmp_llh;
    double *t = cur; cur = tmp; tmp = t;
  }
  memcpy(rates, cur, sizeof a);
  return llh;
}

static double
estimate_rates_collapsed(uint64_t n_nodes, double tol, const double *J,
                         const double *D, const double *input, double *rates) {
  double Jc[8] = {0}, Dc[8] = {0};
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t i = 0; i < 8; ++i) { Jc[i] += J[b*8+i]; Dc[i] += D[b*8+i]; }
  return estimate_rates8(tol, Jc, Dc, input, rates);
}

/* :337-353 */
double
orc_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                   double param_tol, double *rates, double *T, double *baseline) {
  double upd[8], Q[4];
  const double llh = estimate_rates_collapsed(n_nodes, param_tol, J, D, rates, upd);
  memcpy(rates, upd, sizeof upd);
  orc_model_rebuild(rates, T, Q, baseline);
  return llh;
}

/* a16: :318-334 */
void
orc_set_one_change(double *rates, uint64_t n_nodes, double *branches) {
  const double f = orc_rate_scaling_factor(rates);
  for (size_t b = 0; b < n_nodes; ++b) branches[b] = branches[b]*f;
  for (size_t i = 0; i < 8; ++i) rates[i] = rates[i]/f;
}

/* a16: :224-240, :383-422 */
double
orc_estimate_rates_and_branches(uint64_t n_nodes, const double *J, const double *D,
                                double param_tol, double *rates, double *branches,
                                double *T, double *baseline) {
  double upd[8], Q[4];
  estimate_rates_collapsed(n_nodes, param_tol, J, D, rates, upd);
  double *scale = (double *)malloc(n_nodes*sizeof(double));
  double *nb = (double *)malloc(n_nodes*sizeof(double));
  scale[0] = 1.0;
  for (size_t b = 1; b < n_nodes; ++b) {
    double num = 0.0, den = 0.0;
    for (size_t i = 0; i < 8; ++i) num += J[b*8+i];
    for (size_t i = 0; i < 8; ++i) den += D[b*8+i]*upd[i];
    scale[b] = num/den;
  }
  for (size_t b = 0; b < n_nodes; ++b) nb[b] = scale[b]*branches[b];
  orc_set_one_change(upd, n_nodes, nb);
  orc_model_rebuild(upd, T, Q, baseline);
  double Jc[8] = {0}, Dc[8] = {0};
  for (size_t b = 1; b < n_nodes; ++b)
    for (size_t i = 0; i < 8; ++i) {
      Jc[i] += J[b*8+i];
      Dc[i] += scale[b]*D[b*8+i];
    }
  memcpy(rates, upd, sizeof upd);
  memcpy(branches, nb, n_nodes*sizeof(double));
  free(scale); free(nb);
  return orc_log_likelihood(Jc, Dc, rates);
}

/* a17: ParamEstimation.cpp:369-380 */
void
orc_scale_jump_times(uint64_t n_nodes, const double *new_branches) {
  for (size_t s = 0; s < g_n_sites; ++s)
    for (size_t b = 1; b < n_nodes; ++b) {
      path_t *p = &PATH(s, b);
      const double scale = new_branches[b]/p->tot_time;
      for (size_t j = 0; j < p->jumps.n; ++j) p->jumps.v[j] *= scale;
      p->tot_time = new_branches[b];
    }
}
