// TEST INFRASTRUCTURE ONLY — see ../glm.hpp.
#pragma once
#include "../glm.hpp"
