/* TEST INFRASTRUCTURE: see ../core/core.hpp */
#include "../core/core.hpp"
