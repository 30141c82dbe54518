// pmc_oracle_ext.hpp -- oracle part 2: condensation and transport declarations.
// TEST INFRASTRUCTURE ONLY (see pmc_oracle.hpp).
#pragma once
#include "pmc_oracle.hpp"

namespace pmco {
struct TransDomain {};  // filled in by the transport restatement
}  // namespace pmco
