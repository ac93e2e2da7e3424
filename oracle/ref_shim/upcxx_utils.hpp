// TEST INFRASTRUCTURE (oracle/) -- umbrella header + the few string helpers the kcount sources call.
#pragma once
#include <iomanip>
#include <sstream>
#include <string>
#include "upcxx_utils/log.hpp"
#include "upcxx_utils/mem_profile.hpp"
#include "upcxx_utils/progress_bar.hpp"
#include "upcxx_utils/three_tier_aggr_store.hpp"
#include "upcxx_utils/timers.hpp"

