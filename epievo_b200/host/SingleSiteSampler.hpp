/* SingleSiteSampler.hpp -- B200 drop-in for the reference's header of the same name
 * (smithlabcode/epievo src/libepievo/SingleSiteSampler.hpp:35-81).
 *
 * Put this directory BEFORE the reference's src/libepievo on the include path and link
 * SingleSiteSampler_b200.o + libepievo_b200.so instead of the reference's
 * SingleSiteSampler.o: the reference's programs (prog/epievo_est_params_histories.cpp,
 * prog/epievo_sim_pairwise.cpp, harnesses/MCMC_test.cpp) then compile UNCHANGED and run
 * the MCMC sweep on the GPU through the C-ABI in include/epievo_b200.h.  Same class
 * name, same public members, same argument meaning, same error convention
 * (std::runtime_error).  Path / EpiEvoModel / TreeHelper are the reference's own types.
 *
 * Differences a maintainer should know:
 *  - randomness: the caller's std::mt19937 only seeds the device's counter-based Philox
 *    streams (two 32-bit draws per run_mcmc / sweep), so -s still makes runs
 *    reproducible, but chains differ draw-by-draw from the CPU reference (same
 *    stationary law; tests/test_gpu_sampler_stats.py);
 *  - scan order: sites are updated in three stride-3 colour classes per sweep instead
 *    of 1..n-2 sequentially;
 *  - Metropolis_Hastings_site(site_id): the GPU updates whole sweeps.  The call with
 *    site_id == 1 runs one full sweep over sites 1..n-2, calls with other site ids do no
 *    work, so the caller's `for site_id in 1..n-2` loop
 *    (prog/epievo_sim_pairwise.cpp:268-270) performs exactly one sweep, as before.  The
 *    return values of one such loop add up to the sweep's number of accepted updates
 *    (true for site ids 1..n_accepted), not site by site;
 *  - SAMPLE_ROOT must stay false (its value in every caller, SingleSiteSampler.cpp:441):
 *    the device sampler's acceptance ratio relies on the root state being kept; run_mcmc /
 *    Metropolis_Hastings_site throw std::runtime_error otherwise;
 *  - EPIEVO_B200_GPUS=N shards the sites over N GPUs of the box (contiguous blocks, halo
 *    records over NVLink, include/epievo_b200.h epv_multi_*); EPIEVO_B200_DEVICE=k picks
 *    the (first) device.  The chain does not depend on N.
 */
#ifndef SINGLE_SITE_SAMPLER_HPP
#define SINGLE_SITE_SAMPLER_HPP

#include <cstddef>
#include <random>
#include <vector>

#include "Path.hpp"
#include "EpiEvoModel.hpp"
#include "TreeHelper.hpp"

struct epv_multi;

class SingleSiteSampler {
public:
  SingleSiteSampler(const size_t n_burn_in, const size_t n_batch);
  ~SingleSiteSampler();
  SingleSiteSampler(const SingleSiteSampler &) = delete;
  SingleSiteSampler &operator=(const SingleSiteSampler &) = delete;

  void
  reset(const EpiEvoModel &the_model, std::vector<std::vector<Path> > &paths);

  bool
  Metropolis_Hastings_site(const EpiEvoModel &the_model, const TreeHelper &th,
                           const size_t site_id,
                           std::vector<std::vector<Path> > &paths,
                           std::mt19937 &gen);

  void
  run_mcmc(const EpiEvoModel &the_model, const TreeHelper &th,
           std::vector<std::vector<Path> > &paths,
           std::mt19937 &gen,
           std::vector<std::vector<double> > &J_all_sites,
           std::vector<std::vector<double> > &D_all_sites,
           double &acceptance_rate);

  // MCMC parameter constants (same public fields as the reference)
  bool SAMPLE_ROOT;
  size_t burn_in;
  size_t batch;

private:
  void configure(const EpiEvoModel &the_model, const TreeHelper &th);
  void upload(const std::vector<std::vector<Path> > &paths, const TreeHelper &th);
  void download(std::vector<std::vector<Path> > &paths);
  void check(int rc) const;

  epv_multi *ctx;          // one handle over EPIEVO_B200_GPUS contexts (default 1)
  size_t sweep_counter;
  size_t last_sweep_accepts;
  // flat exchange buffers, reused across calls
  std::vector<unsigned char> init_state;
  std::vector<unsigned long long> row_off;
  std::vector<double> jumps, tot_time;
};

/* GPU versions of the two whole-data passes of ParamEstimation.hpp:63-80 that the
 * estimators call on paths (a one-line change in prog/epievo_est_complete.cpp, see
 * INTEGRATION.md); same signatures as the reference's functions. */
namespace epievo_b200 {
void
get_sufficient_statistics(const std::vector<std::vector<Path> > &all_paths,
                          const TreeHelper &th,
                          std::vector<std::vector<double> > &J,
                          std::vector<std::vector<double> > &D);
/* The reference's own signature (no tree: J/D are per branch and do not depend on the topology);
 * host/ParamEstimation_b200.cpp makes it the link-time replacement of the reference's function. */
void
get_sufficient_statistics(const std::vector<std::vector<Path> > &all_paths,
                          std::vector<std::vector<double> > &J,
                          std::vector<std::vector<double> > &D);
}

#endif
