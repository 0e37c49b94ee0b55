// TEST INFRASTRUCTURE ONLY: include-path forwarder so that the unmodified reference sources find "<opencv2/opencv.hpp>".
#pragma once
#include "../cvmat_shim.hpp"
