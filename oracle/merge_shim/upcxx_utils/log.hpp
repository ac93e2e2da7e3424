// TEST INFRASTRUCTURE (oracle/) -- the logging macros of oracle/ref_shim, over this directory's upcxx stand-in.
#pragma once
#include "../../ref_shim/upcxx_utils/log.hpp"
