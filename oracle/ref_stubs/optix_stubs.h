// TEST INFRASTRUCTURE ONLY — see optix.h.
#pragma once
