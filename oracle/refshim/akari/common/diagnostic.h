// Shadows src/akari/common/diagnostic.h:12 (`#pragma GCC diagnostic error "-Wall"`), which turns
// gcc-13 sign-compare warnings in kernel/bvh-accelerator.h:198,199,496 into hard errors.
// Semantics-preserving; oracle/_ref build only.
#pragma once
