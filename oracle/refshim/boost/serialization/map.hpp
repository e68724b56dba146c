// Stand-in, see serialization.hpp (TEST INFRASTRUCTURE).
#pragma once
#include "serialization.hpp"
