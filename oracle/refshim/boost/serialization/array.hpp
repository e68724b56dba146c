// Stand-in for <boost/serialization/array.hpp>: the reference's MapPoint.h includes it, nothing on the compiled path uses it.
#pragma once
#include "serialization.hpp"
