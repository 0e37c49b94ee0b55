// shim forwarding header (test infrastructure): see cvshim.hpp
#pragma once
#include <iterator>
#include <list>
#include "../../cvshim.hpp"
