/* ParamEstimation_b200.cpp -- link-time drop-ins for the two functions of the reference's
 * ParamEstimation.hpp that walk every path on the host:
 *
 *   get_sufficient_statistics(all_paths, J, D)            per branch, ParamEstimation.cpp:92-114
 *   estimate_rates(VERBOSE, param_tol, all_paths, model)  ParamEstimation.cpp:355-366
 *
 * Same signatures and results (J exact; D to the 2^-40 fixed-point quantum of the device pass, 1e-12
 * relative), computed by epv_suffstats on the GPU.  The M-step is the reference's own
 * estimate_rates(VERBOSE, param_tol, J, D, model) from its ParamEstimation.o.  Link this object BEFORE
 * the reference's ParamEstimation.o with -Wl,--allow-multiple-definition (the first definition wins):
 * prog/epievo_est_complete.cpp then runs on the GPU with no source change
 * (integration/_build/epievo_est_complete_b200).  The `-b` overload
 * estimate_rates_and_branches(VERBOSE, tol, all_paths, th, model) (ParamEstimation.cpp:426-446) indexes
 * its argument as [site][node] in one loop and [node][site] in another; it is left to the reference. */
#include <iostream>
#include <vector>

#include "Path.hpp"
#include "EpiEvoModel.hpp"
#include "TreeHelper.hpp"
#include "ParamEstimation.hpp"
#include "SingleSiteSampler.hpp"          // epievo_b200::get_sufficient_statistics (drop-in header)

using std::vector;

void
get_sufficient_statistics(const vector<vector<Path> > &all_paths,
                          vector<vector<double> > &J, vector<vector<double> > &D) {
  epievo_b200::get_sufficient_statistics(all_paths, J, D);
}

double
estimate_rates(const bool VERBOSE, const double param_tol,
               const vector<vector<Path> > &all_paths, EpiEvoModel &the_model) {
  if (VERBOSE)
    std::cerr << "[COMPUTING INITIAL SUFFICIENT STATISTICS]" << std::endl;
  vector<vector<double> > J, D;
  epievo_b200::get_sufficient_statistics(all_paths, J, D);
  return estimate_rates(VERBOSE, param_tol, J, D, the_model);
}
