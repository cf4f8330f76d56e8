// Force-included (-include) ahead of every reference translation unit when the
// oracle builds /root/reference sources in place.  Test infrastructure only.
// Mirrors what the engine's precompiled header provides (cave/engine/pch.h).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <utility>

#include "engine/pch.h"
