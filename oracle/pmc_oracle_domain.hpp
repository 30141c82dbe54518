// pmc_oracle_domain.hpp -- the object behind the oracle's C handle.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
#pragma once
#include "pmc_oracle.hpp"
#include "pmc_oracle_ext.hpp"
#include <memory>
#include <string>
#include <vector>

using namespace pmco;

struct OracleDomain {
  AeroData data;
  std::vector<AeroState> cells;
  std::vector<EnvState> env;
  std::unique_ptr<Rng> rng;
  std::vector<int64_t> global_cell;  // global cell index of each local cell (rng address + id rank)
  std::string last_error;
  int cond_method = 0;  // 0: CondenseOpts default; set by pmco_set_condense_method
  std::vector<double> cond_h;  // per cell: step size carried between condense calls (0 = none yet)
  TransDomain trans;  // transport geometry + met + probabilities (pmc_oracle_ext.hpp)
};

#define ORACLE_TRY try {
#define ORACLE_CATCH(D)                       \
  }                                           \
  catch (const PmcError& e) {                 \
    (D)->last_error = e.what();               \
    return e.code ? e.code : -1;              \
  }                                           \
  catch (const std::exception& e) {           \
    (D)->last_error = e.what();               \
    return -1;                                \
  }                                           \
  return 0;

