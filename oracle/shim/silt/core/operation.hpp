// oracle/shim/silt/core/operation.hpp - TEST INFRASTRUCTURE ONLY.  path.cu includes it; nothing of it is used there.
#pragma once
