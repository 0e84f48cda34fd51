/* epievo_b200.h -- C-ABI of the B200-native MCMC / sufficient-statistics hot path.
 *
 * Drop-in boundary for the reference's (smithlabcode/epievo @ 6e0525a) C++ API of
 * this path.  The reference has no FFI layer; what its three programs and one
 * harness call directly is listed beside each entry point ("replaces ..."), with
 * paths relative to the reference's src/libepievo/.  The C++ shim that keeps the
 * reference's class and function signatures on top of these calls is
 * epievo_b200/host/SingleSiteSampler.hpp (+ IndepSite_b200.cpp); INTEGRATION.md shows how a
 * maintainer links it.  Precondition inherited from the reference's callers: SAMPLE_ROOT is false
 * (SingleSiteSampler.cpp:441; the sampler's acceptance ratio relies on the root state being kept).
 *
 * Conventions
 *  - every call returns 0 on success, non-zero on failure; epv_last_error() then
 *    holds the message (the shim rethrows it as std::runtime_error, the reference's
 *    error convention, e.g. prog/epievo_est_params_histories.cpp:296-299);
 *  - buffers passed in are caller-owned and only read during the call; output
 *    buffers are caller-allocated;
 *  - one host thread drives one context; one context drives one GPU;
 *  - there is NO CPU fallback: without a CUDA device epv_create fails.
 *
 * Path exchange format ("node-major CSR"), mirroring vector<vector<Path>>
 * (Path.hpp:33-57) as transposed by the callers
 * (prog/epievo_est_params_histories.cpp:180-192):
 *    init_state : uint8  [n_nodes * n_sites]       row = node*n_sites + site
 *    row_off    : uint64 [n_nodes * n_sites + 1]   offsets into jumps
 *    jumps      : double [row_off[n_nodes*n_sites]]
 *    tot_time   : double [n_nodes]                 Path::tot_time of the node's rows
 * Node 0 (root) rows are empty; the root state of site s is init_state of node 1
 * (SingleSiteSampler.cpp:409-410, 502-504).
 */
#ifndef EPIEVO_B200_H
#define EPIEVO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct epv_ctx epv_ctx;

#define EPV_N_TRIPLETS 8

/* ---- context ------------------------------------------------------------------ */
/* device: CUDA device ordinal.  Fails (no fallback) when CUDA is unavailable. */
int  epv_create(epv_ctx **ctx, int device);
void epv_destroy(epv_ctx *ctx);
const char *epv_last_error(const epv_ctx *ctx);
/* Run all subsequent work on this cudaStream_t (0 = the context's own stream). */
int  epv_set_stream(epv_ctx *ctx, void *cuda_stream);
int  epv_synchronize(epv_ctx *ctx);

/* ---- inputs ---------------------------------------------------------------------- */
/* replaces TreeHelper (TreeHelper.hpp:38-57): preorder arrays, branch b = edge above
 * node b, branches[0] ignored. */
int epv_set_tree(epv_ctx *ctx, uint64_t n_nodes, const uint64_t *subtree_sizes,
                 const uint64_t *parent_ids, const double *branches);
/* replaces the EpiEvoModel fields the path reads (EpiEvoModel.hpp:32-65):
 * triplet_rates[8] (index l*4+m*2+r) and the horizontal 2x2 T (row-major). */
int epv_set_model(epv_ctx *ctx, const double *triplet_rates, const double *T);

/* Sharding (one context per GPU): this context holds global sites
 * [site_begin - halo_lo, site_end + halo_hi) of an n_global-site problem; only sites
 * in [site_begin, site_end) are updated / counted here.  site_begin must be a
 * multiple of 3; halo_lo / halo_hi are 3 on a side with a neighbour, 0 at a global
 * end; a shard owns at least 6 sites.  Default (never called): the whole problem. */
int epv_set_shard(epv_ctx *ctx, uint64_t n_global, uint64_t site_begin,
                  uint64_t site_end, uint64_t halo_lo, uint64_t halo_hi);

/* host paths -> device CSR (replaces handing vector<vector<Path>>& to the sampler).
 * n_sites counts the rows actually passed (shard + halos). */
int epv_upload_paths(epv_ctx *ctx, uint64_t n_sites, uint64_t n_nodes,
                     const uint8_t *init_state, const uint64_t *row_off,
                     const double *jumps, const double *tot_time);
/* device CSR -> host.  Call epv_paths_size first to size the buffers. */
int epv_paths_size(epv_ctx *ctx, uint64_t *n_sites, uint64_t *n_nodes,
                   uint64_t *n_jumps);
int epv_download_paths(epv_ctx *ctx, uint8_t *init_state, uint64_t *row_off,
                       double *jumps, double *tot_time);
/* The same exchange at 2 bytes per row instead of 9 (the per-row arrays dominate the PCIe
 * traffic: a row of Path.hpp:38-48 holds 0.1-2 jumps): row_code[node*n_sites + site] =
 * number of jumps of the row | init_state << 15 (root rows: count 0); jumps as above, in row
 * order; n_jumps = their number (checked against the codes).  Offsets and state bit sets
 * are built on the device.  A row with more than 32767 jumps needs the format above. */
int epv_upload_paths_compact(epv_ctx *ctx, uint64_t n_sites, uint64_t n_nodes,
                             const uint16_t *row_code, const double *jumps,
                             uint64_t n_jumps, const double *tot_time);
/* jumps holds room for jumps_cap values; *n_jumps receives the number needed/written. */
int epv_download_paths_compact(epv_ctx *ctx, uint16_t *row_code, double *jumps,
                               uint64_t jumps_cap, uint64_t *n_jumps, double *tot_time);

/* ---- the sampler (SingleSiteSampler.hpp:35-81) --------------------------------- */
/* replaces SingleSiteSampler::reset (SingleSiteSampler.cpp:449-475): caches the
 * per-site triplet log-likelihoods for the current model and paths. */
int epv_reset(epv_ctx *ctx);
/* replaces SingleSiteSampler::single_iteration (SingleSiteSampler.cpp:538-548):
 * one Metropolis-Hastings update of every interior site, applied in three stride-3
 * colour passes; randomness is Philox4x32-10 keyed by (seed, sweep_idx, site, node).
 * n_accepted may be NULL. */
int epv_sweep(epv_ctx *ctx, uint64_t seed, uint64_t sweep_idx, uint64_t *n_accepted);
/* replaces SingleSiteSampler::run_mcmc (SingleSiteSampler.cpp:550-598): burn_in
 * sweeps, then batch x (sweep + per-branch J/D), batch-averaged.  J, D are
 * [n_nodes*8] (row 0 zero); acc_rate = accepted / (batch*(n_global-2)) over this
 * shard's sites.  Sweep indices first_sweep .. first_sweep+burn_in+batch-1. */
int epv_run_mcmc(epv_ctx *ctx, uint64_t burn_in, uint64_t batch, uint64_t seed,
                 uint64_t first_sweep, double *J, double *D, double *acc_rate);
/* Same as epv_run_mcmc but leaves J/D sums (not averaged) and the accept count on
 * the device-side accumulators and returns them raw, so a multi-GPU driver can
 * all-reduce them before averaging. */
int epv_run_mcmc_raw(epv_ctx *ctx, uint64_t burn_in, uint64_t batch, uint64_t seed,
                     uint64_t first_sweep, double *J_sum, double *D_sum,
                     uint64_t *n_accepted);

/* ---- sufficient statistics, likelihood, gradient -------------------------------- */
/* replaces get_sufficient_statistics(paths, J, D) per branch
 * (ParamEstimation.cpp:92-114): J, D are [n_nodes*8], row 0 zero.  J is exact; D is summed in
 * 64-bit fixed point (quantum tot_time * 2^-40 per term, far inside the 1e-9 parity bound),
 * which makes the result independent of the order in which thread blocks finish: two contexts
 * holding the same paths return bit-identical D. */
int epv_suffstats(epv_ctx *ctx, double *J, double *D);
/* replaces log_likelihood + get_gradient (ParamEstimation.cpp:131-194) evaluated on
 * the device-resident J/D of the last epv_suffstats / epv_run_mcmc, collapsed over
 * branches. */
int epv_llh_grad(epv_ctx *ctx, const double *rates, double *llh, double *grad);
/* replaces scale_jump_times (ParamEstimation.cpp:369-380): every jump of branch b
 * is multiplied by new_branches[b]/tot_time[b]; tot_time and the tree's branch
 * lengths become new_branches. */
int epv_scale_jump_times(epv_ctx *ctx, const double *new_branches);

/* One whole MCEM iteration with the paths RESIDENT on the device -- replaces the body of the
 * loop at prog/epievo_est_params_histories.cpp:239-264: reset, run_mcmc (E-step), the M-step
 * (estimate_rates + set_one_change_per_site_per_unit_time, or estimate_rates_and_branches
 * when optimize_branches != 0) and scale_jump_times.  rates[8], T[4] and branches[n_nodes]
 * are the current model / tree on entry and the updated ones on return; baseline[4]
 * receives the stationary baseline (may be NULL).  Paths never leave the GPU; call
 * epv_download_paths when the reference would write its local-paths file. */
int epv_mcem_iteration(epv_ctx *ctx, uint64_t burn_in, uint64_t batch, uint64_t seed,
                       uint64_t first_sweep, int optimize_branches, double param_tol,
                       double *rates, double *T, double *branches, double *baseline,
                       double *acc_rate, double *llh);

/* ---- multi-GPU building blocks (see epv_set_shard) --------------------------------- */
/* One stride-3 colour pass (colour 0,1,2) of sweep sweep_idx over this shard's sites;
 * epv_sweep is three of these.  A sharded driver exchanges halos between passes. */
int epv_pass(epv_ctx *ctx, int colour, uint64_t seed, uint64_t sweep_idx);
/* Accept counter of the passes issued since the last call with reset != 0. */
int epv_accept_count(epv_ctx *ctx, int reset, uint64_t *n_accepted);

/* Halo exchange.  After colour pass c, exactly one of the three boundary sites on each
 * side of a shard can have changed (radius-2 dependence, stride-3 colours).
 * epv_halo_pack writes that site's record -- rows of every branch, state bits, the three
 * tri_llh entries its update wrote, its global index -- to the DEVICE buffers rec_lo
 * (for the lower-index neighbour) and rec_hi (for the higher-index neighbour), each
 * epv_halo_record_bytes long; the driver moves them to the neighbours (NCCL send/recv or
 * a peer copy) and calls epv_halo_unpack with the records received FROM the lower
 * (rec_lo) and higher (rec_hi) neighbour.  Pass NULL for a side that has no neighbour.
 * All work is enqueued on the context's stream. */
int epv_halo_record_bytes(epv_ctx *ctx, uint64_t *bytes);
int epv_halo_pack(epv_ctx *ctx, int colour, void *rec_lo, void *rec_hi);
int epv_halo_unpack(epv_ctx *ctx, int colour, const void *rec_lo, const void *rec_hi);
/* A halo row (and a record) holds up to `capacity` jumps per branch: sized from the rows next to the
 * shard's ends at upload (at least 48, twice the largest), the same on every shard of a run -- the
 * driver takes the maximum over shards and sets it everywhere.  *needed is what the boundary rows
 * packed so far ask for (>= capacity when they passed half of it): call between sweeps and grow
 * (contents are kept; epv_halo_record_bytes changes).  A row that still outgrows a record inside one
 * sweep raises an error, it is never truncated. */
int epv_halo_capacity(epv_ctx *ctx, uint64_t *capacity, uint64_t *needed);
int epv_halo_set_capacity(epv_ctx *ctx, uint64_t capacity);
/* The same check without a host synchronisation (one per sweep costs a multi-process driver ~0.3 ms of idle
 * GPU: 0.4 ms x 20 sweeps was the whole weak-scaling loss at 8 GPUs): enqueue, on `stream` (a cudaStream_t the
 * caller has ordered after the packs; NULL = the context's side stream), a copy of the longest boundary row
 * packed so far into *dst_device (one uint64, device memory).  The driver reduces it over shards (max), reads
 * it a sweep later and grows with epv_halo_set_capacity when twice the value passes the capacity. */
int epv_halo_need_async(epv_ctx *ctx, void *dst_device, void *stream);

/* The same colour pass with the exchange OVERLAPPED.  epv_pass_begin enqueues, on the context's side
 * stream (epv_set_side_stream; default: an internal one), the halo-aware update of the few groups next
 * to the shard's ends and the pack of the boundary sites' records (taken from the proposal staging,
 * before the CSR is rebuilt) into rec_lo / rec_hi, and on the main stream the interior update and the
 * CSR rebuild.  The driver moves the records ON THE SIDE STREAM (peer copy, NCCL send/recv) and calls
 * epv_pass_end with the records received, which unpacks them on the side stream and makes the main
 * stream's next work wait for that; after colour 2 it also does the per-sweep housekeeping (one host
 * synchronisation).  Same chain, bit for bit, as epv_pass + epv_halo_pack/unpack. */
int epv_set_side_stream(epv_ctx *ctx, void *cuda_stream);
int epv_pass_begin(epv_ctx *ctx, int colour, uint64_t seed, uint64_t sweep_idx, void *rec_lo, void *rec_hi);
int epv_pass_end(epv_ctx *ctx, int colour, const void *rec_lo, const void *rec_hi);
/* J/D statistics as exact integer accumulators, for drivers of several shards (the fold of
 * epv_suffstats / epv_run_mcmc is a function of integer sums only): epv_stats_begin clears the context's
 * accumulators, epv_stats_accumulate adds one J/D pass over the current paths (enqueued only: call it
 * after every batch sweep), epv_stats_export copies them -- epv_stats_words 8-byte words -- to a DEVICE
 * buffer of the caller, which sums the shards' buffers element-wise as 64-bit integers (an all-reduce
 * on int64, or peer copies + epv_stats_add), and epv_stats_fold turns a summed buffer (NULL: the
 * context's own accumulators) into J / D [n_nodes*8] on the host.  J and D then do not depend on how the
 * sites were sharded, to the last bit. */
int epv_stats_words(epv_ctx *ctx, uint64_t *words);
int epv_stats_begin(epv_ctx *ctx);
int epv_stats_accumulate(epv_ctx *ctx);
int epv_stats_export(epv_ctx *ctx, void *dev_words);
int epv_stats_add(epv_ctx *ctx, const void *dev_words);
int epv_stats_fold(epv_ctx *ctx, const void *dev_words, double *J, double *D);
/* epv_suffstats WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernels and same result as epv_suffstats on the context's resident
 * paths (J exact, D bit-identical: integer accumulation); the scratch the kernels accumulate into is a
 * file-scope device array of the library (csrc/api_noalloc.cu) instead of the context's buffer.
 * LIMIT: trees of at most 64 nodes (the array is 40 320 bytes; every GPU that loads the library holds it
 * until the process exits; epv_static_limits reports both numbers).  A larger tree is rejected with an
 * error, nothing is launched.  The call itself calls no cudaMalloc / cudaFree / allocating routine.
 * ONE array per GPU, shared by every context and every call: each call clears it first, and calls that
 * overlap on the same GPU (other contexts / streams / host threads) race on it -- serialise them. */
int epv_suffstats_static(epv_ctx *ctx, double *J, double *D);
int epv_static_limits(uint64_t *max_nodes, uint64_t *scratch_bytes);
/* epv_suffstats in two halves: enqueue the pass / read the result (so that a driver of several
 * contexts has all of them running before it waits for the first). */
int epv_suffstats_async(epv_ctx *ctx);
int epv_suffstats_fetch(epv_ctx *ctx, double *J, double *D);

/* ---- several GPUs behind one handle, one host process (csrc/api_multi.cu) --------------------------
 * What the C++ drop-in uses when EPIEVO_B200_GPUS=N: the calls above with the same meaning and the
 * same exchange format, over n_shards contexts holding contiguous site blocks (starts are multiples of
 * 3), halo records moved by peer copies (NVLink) on the side streams, J/D sums added on the host.
 * devices: n_shards CUDA ordinals, or NULL for 0..n_shards-1 (ordinals may repeat: several shards on
 * one GPU, used by the tests).  n_shards = 1 is the single context.  The chain does not depend on the
 * number of shards (Philox is keyed by the global site index). */
typedef struct epv_multi epv_multi;
int  epv_multi_create(epv_multi **m, int n_shards, const int *devices);
void epv_multi_destroy(epv_multi *m);
const char *epv_multi_last_error(const epv_multi *m);
int  epv_multi_shards(const epv_multi *m);
epv_ctx *epv_multi_context(epv_multi *m, int shard);
int epv_multi_partition(uint64_t n_global, int n_shards, uint64_t *begin, uint64_t *end);
int epv_multi_set_tree(epv_multi *m, uint64_t n_nodes, const uint64_t *subtree_sizes,
                       const uint64_t *parent_ids, const double *branches);
int epv_multi_set_model(epv_multi *m, const double *triplet_rates, const double *T);
int epv_multi_upload_paths(epv_multi *m, uint64_t n_sites, uint64_t n_nodes, const uint8_t *init_state,
                           const uint64_t *row_off, const double *jumps, const double *tot_time);
int epv_multi_paths_size(epv_multi *m, uint64_t *n_sites, uint64_t *n_nodes, uint64_t *n_jumps);
int epv_multi_download_paths(epv_multi *m, uint8_t *init_state, uint64_t *row_off, double *jumps,
                             double *tot_time);
int epv_multi_reset(epv_multi *m);
int epv_multi_sweep(epv_multi *m, uint64_t seed, uint64_t sweep_idx, uint64_t *n_accepted);
int epv_multi_run_mcmc(epv_multi *m, uint64_t burn_in, uint64_t batch, uint64_t seed, uint64_t first_sweep,
                       double *J, double *D, double *acc_rate);
int epv_multi_suffstats(epv_multi *m, double *J, double *D);
int epv_multi_scale_jump_times(epv_multi *m, const double *new_branches);
int epv_multi_mcem_iteration(epv_multi *m, uint64_t burn_in, uint64_t batch, uint64_t seed,
                             uint64_t first_sweep, int optimize_branches, double param_tol, double *rates,
                             double *T, double *branches, double *baseline, double *acc_rate, double *llh);
uint64_t epv_multi_launch_count(const epv_multi *m);

/* ---- deterministic per-site intermediates (parity tests of a2/a3/a4/a7) --------- */
/* For local site `site`: the neighbour-interval lists of every node
 * (collect_segment_info, Segment.cpp:35-79), the Felsenstein p/q values (pruning,
 * SingleSiteSampler.cpp:80-157), the root posterior (:167-176) and the proposal
 * log-probability of the CURRENT path (proposal_prob, :316-339), computed by the
 * same device functions the sampler kernel uses.  Returns the interval count in
 * *n_intervals; fails if it exceeds cap. */
int epv_site_debug(epv_ctx *ctx, uint64_t site, uint64_t cap, uint64_t *n_intervals,
                   uint64_t *seg_off, double *rate0, double *rate1, uint64_t *trip0,
                   uint64_t *trip1, double *len, double *p, double *q,
                   double *root_p0, double *proposal_logp);
/* Cached tri_llh of every local site (SingleSiteSampler.cpp:470-474). */
int epv_get_tri_llh(epv_ctx *ctx, double *tri_llh);
/* Draw n end-conditioned paths on one interval with the sampler kernel's own device
 * routine (replaces end_cond_sample_forward_rejection, EndCondSampling.cpp:479-509;
 * same law, rejection-free): writes the jump count of each draw to n_jumps[n] and
 * the first `cap` jump times of each to jumps[n*cap]. */
int epv_end_cond_sample(epv_ctx *ctx, double rate0, double rate1, int start_state,
                        int end_state, double T, uint64_t seed, uint64_t n,
                        uint64_t cap, uint32_t *n_jumps, double *jumps);

/* ---- host-side model arithmetic (O(8); no device work) ---------------------------- */
/* replaces read_model/initialize/scale_triplet_rates (EpiEvoModel.cpp:319-417). */
void epv_model_from_params(double T00, double T11, double baseline00,
                           double baseline11, int scale, double *rates, double *T,
                           double *Q, double *baseline);
/* replaces EpiEvoModel::rebuild_from_triplet_rates (EpiEvoModel.cpp:420-449). */
void epv_model_rebuild(const double *rates, double *T, double *Q, double *baseline);
/* replaces rate_scaling_factor (EpiEvoModel.cpp:158-189). */
double epv_rate_scaling_factor(const double *rates);
/* replaces estimate_rates(J, D, model) (ParamEstimation.cpp:337-353): rates in/out. */
double epv_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                          double param_tol, double *rates, double *T,
                          double *baseline);
/* replaces estimate_rates_and_branches(J, D, th, model) (:383-422). */
double epv_estimate_rates_and_branches(uint64_t n_nodes, const double *J,
                                       const double *D, double param_tol,
                                       double *rates, double *branches, double *T,
                                       double *baseline);
/* replaces set_one_change_per_site_per_unit_time (:318-334). */
void epv_set_one_change(double *rates, uint64_t n_nodes, double *branches);

/* ---- site-independent E-step of epievo_initialization -------------- */
/* Every site evolves as one 2-state CTMC with rates[2] = (0->1, 1->0) on every branch
 * (IndepSite.cpp).  J and D are [n_nodes][2] (per start state; node 0 unused) and hold SUMS
 * over the context's owned sites: compute_sufficient_statistics divides them by n_sites
 * (IndepSite.cpp:296-301), expectation_sufficient_statistics does not (:228-238).
 * A sharded context works on its owned sites; after epv_indep_update_paths refresh the
 * halos (epv_halo_pack / unpack for colours 0, 1, 2) before the next MH pass. */
/* replaces compute_sufficient_statistics (IndepSite.cpp:266-303), as sums. */
int epv_indep_suffstats(epv_ctx *ctx, double *J, double *D);
/* replaces expectation_sufficient_statistics (IndepSite.cpp:221-239). */
int epv_indep_expectation(epv_ctx *ctx, const double *rates, double *J, double *D);
/* replaces update_paths_indep (IndepSite.cpp:243-260): every owned site gets a new history
 * given its leaf states and root state; Philox streams keyed (seed; iteration, site, node). */
int epv_indep_update_paths(epv_ctx *ctx, const double *rates, uint64_t seed,
                           uint64_t iteration);
/* host: per-branch tables (continuous_time_trans_prob_mat, expectation_J, expectation_D,
 * ContinuousTimeMarkovModel.cpp:142-226), 2x2 row-major over (start, end). */
void epv_indep_tables(double r0, double r1, double T, double *P, double *J0, double *J1,
                      double *D0, double *D1);
/* host: estimate_rates_indep (IndepSite.cpp:306-326). */
void epv_indep_estimate_rates(uint64_t n_nodes, const double *J, const double *D,
                              double *rates);
/* host: estimate_rates_and_branches_indep (:341-374) up to the path rescaling, which is
 * epv_scale_jump_times(ctx, branches). */
void epv_indep_estimate_rates_and_branches(uint64_t n_nodes, const double *J,
                                           const double *D, double *rates,
                                           double *branches);

/* ---- forward simulation of histories --------------------------------- */
/* replaces epievo_sim's simulation (prog/epievo_sim.cpp:102-152 sample_jump, :326-356 the loop
 * over nodes; TripletSampler.cpp) followed by global_jumps_to_paths: evolves a root sequence down
 * the tree set with epv_set_tree under the triplet rates set with epv_set_model, every interior
 * site flipping at rate[left, self, right], the two end sites fixed.  Same law as the reference's
 * global Gillespie simulation (thinning of per-site candidate events, decided in time order
 * between neighbours), not the same random stream.  root_states: n_sites bytes (0/1), or NULL to
 * draw the root from the horizontal chain T (sample_state_sequence, EpiEvoModel.cpp:282-298).
 * Philox streams keyed (seed; candidate, site, node): the result depends on seed only.
 * The histories stay on the device; *n_jumps tells the size of the jumps array to fetch. */
int epv_simulate(epv_ctx *ctx, uint64_t n_sites, const uint8_t *root_states, uint64_t seed,
                 uint64_t *n_jumps);
/* copies the result out in the epv_upload_paths exchange format: init_state [n_nodes][n_sites],
 * row_off [n_nodes*n_sites + 1] (root rows empty), jumps [n_jumps]; frees the device copy. */
int epv_simulate_fetch(epv_ctx *ctx, uint8_t *init_state, uint64_t *row_off, double *jumps);
/* or: the simulated histories become the context's paths without leaving the device (what
 * epv_simulate_fetch + epv_upload_paths would load; honours epv_set_shard, n_sites of epv_simulate =
 * the shard's local sites; refresh the halos from the neighbours afterwards). */
int epv_simulate_adopt(epv_ctx *ctx);

/* ---- local-paths text files (host only, multi-threaded) ------------- */
/* replaces read_paths (Path.cpp:123-148): parses the file with n_threads threads
 * (<= 0: all cores) into an opaque handle and reports the sizes; epv_pathfile_get fills
 * caller-allocated arrays in the epv_upload_paths exchange format (root rows empty);
 * node names as in the file.  Errors: non-zero + epv_io_last_error() (thread-local),
 * "cannot read: <file>" as the reference throws. */
typedef struct epv_pathfile epv_pathfile;
int epv_pathfile_read(const char *filename, int n_threads, epv_pathfile **out,
                      uint64_t *n_sites, uint64_t *n_nodes, uint64_t *n_jumps);
int epv_pathfile_get(const epv_pathfile *f, uint8_t *init_state, uint64_t *row_off,
                     double *jumps, double *tot_time);
const char *epv_pathfile_node_name(const epv_pathfile *f, uint64_t node);
void epv_pathfile_free(epv_pathfile *f);
/* replaces write_root_to_pathfile_local + append_to_pathfile_local
 * (prog/epievo_est_params_histories.cpp:56-75) with operator<<(ostream, Path)
 * (Path.cpp:62-71): byte-identical output, formatted by n_threads threads. */
int epv_pathfile_write(const char *filename, uint64_t n_sites, uint64_t n_nodes,
                       const char *const *node_names, const uint8_t *init_state,
                       const uint64_t *row_off, const double *jumps,
                       const double *tot_time, int n_threads);
/* The same content as a binary file (2 bytes per row + 8 per jump instead of ~22 bytes of text per row; the
 * "optional binary CSR sidecar" of SURVEY 8f-1 -- no counterpart in the reference): "EPVPATH1" header with the
 * sizes and a checksum, node names, tot_time, one u16 per row (bit 15 = init_state, bits 0-14 = jumps of the
 * row, hence at most 32767 jumps per row), the jump times.  epv_pathfile_read recognises the magic, so every
 * reader of the text form reads this one too; a truncated or altered file is an error, never partial data.
 * host/read_paths_b200.cpp (the read_paths drop-in) prefers <file>.epvb when it is not older than <file>. */
int epv_pathfile_write_binary(const char *filename, uint64_t n_sites, uint64_t n_nodes,
                              const char *const *node_names, const uint8_t *init_state,
                              const uint64_t *row_off, const double *jumps,
                              const double *tot_time, int n_threads);
const char *epv_io_last_error(void);

/* ---- instrumentation ------------------------------------------------------------- */
/* Kernel launches issued by this context since creation (bench "gpu_launches"). */
uint64_t epv_launch_count(const epv_ctx *ctx);
/* How often a colour pass outgrew a branch's jump storage and was recovered (storage regrown, the CSR
 * rebuild rerun from its intact inputs).  The reference's vector<double> rows are unbounded
 * (Path.hpp:33-57); here capacity is 3 x the fill and is checked after every pass while headroom is thin
 * (a branch more than half full, or holding less than 1.5 x its stationary expectation), else per sweep. */
uint64_t epv_recovery_count(const epv_ctx *ctx);
/* Device time (ms, CUDA events on the context's stream) and launch count of the
 * colour-pass update kernel (mh_update_kernel) accumulated since the last call with
 * reset != 0; epv_rebuild_kernel_time returns the csr_rebuild_kernel time accumulated
 * over the same launches, as of the last epv_mh_kernel_time call. */
int epv_mh_kernel_time(epv_ctx *ctx, int reset, double *ms, uint64_t *launches);
int epv_rebuild_kernel_time(epv_ctx *ctx, double *ms);
int epv_enable_kernel_timing(epv_ctx *ctx, int on);
/* Total jumps currently stored, per node ([n_nodes], entry 0 zero). */
int epv_jump_counts(epv_ctx *ctx, uint64_t *per_node);

#ifdef __cplusplus
}
#endif
#endif
