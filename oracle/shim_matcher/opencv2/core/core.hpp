// TEST INFRASTRUCTURE ONLY: include-path forwarder so that the unmodified reference sources find "<opencv2/core/core.hpp>".
#pragma once
#include "../../cvmat_shim.hpp"
