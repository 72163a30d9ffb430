// TEST INFRASTRUCTURE ONLY — the reference generates this header with cmake (GenerateExportHeader); the host driver
// needs only the empty export macro.
#pragma once
#define RAY_TRACER_FACILITY_API
