#pragma once
#include "../../include/madrona/b200/static_map.hpp"
