// TEST INFRASTRUCTURE ONLY (oracle/): see opencv2/core/mat.hpp in this directory.
#pragma once
#include "mat.hpp"
