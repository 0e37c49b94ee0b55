// TEST INFRASTRUCTURE ONLY: include-path forwarder so that the unmodified reference sources find "<opencv2/features2d/features2d.hpp>".
#pragma once
#include "../../cvmat_shim.hpp"
