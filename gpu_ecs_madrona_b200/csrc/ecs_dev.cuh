// The device side of the ECS state layer is a public header (user systems are compiled
// against it, include/madrona/b200/): this forwarder keeps the library's own includes short.
#pragma once
#include "../../include/madrona_b200_apps.h"
#include "../../include/madrona/b200/ecs_device.cuh"
