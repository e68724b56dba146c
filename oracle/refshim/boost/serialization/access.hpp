// Stand-in: the reference's headers include this boost.serialization header; nothing on the compiled path uses it.
#pragma once
#include "serialization.hpp"
