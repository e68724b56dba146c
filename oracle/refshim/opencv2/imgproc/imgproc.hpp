// Stand-in: see ../opencv.hpp (TEST INFRASTRUCTURE).
#include "../opencv.hpp"
