// Column ids of the rigid-body archetype -> table pointers / element sizes
// (reference: phys::Cols, include/madrona/physics.hpp:449-464; component sizes
// physics.hpp:150-171,402-415, components.hpp:9-29).
#pragma once
#include "phys_types.cuh"
#include "../../include/madrona_b200.h"

namespace mdr {

static const int32_t kColBytes[MDR_NUM_PHYS_COLS] = { 8, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4 };

inline void **colSlot(PhysTables &t, int col)
{
    switch (col) {
    case MDR_COL_ENTITY: return (void **)&t.entity;
    case MDR_COL_POSITION: return (void **)&t.pos;
    case MDR_COL_ROTATION: return (void **)&t.rot;
    case MDR_COL_SCALE: return (void **)&t.scale;
    case MDR_COL_VELOCITY: return (void **)&t.vel;
    case MDR_COL_OBJECT_ID: return (void **)&t.obj;
    case MDR_COL_RESPONSE_TYPE: return (void **)&t.resp;
    case MDR_COL_SUBSTEP_PREV: return (void **)&t.prev;
    case MDR_COL_PRESOLVE_POS: return (void **)&t.prex;
    case MDR_COL_PRESOLVE_VEL: return (void **)&t.prev_vel;
    case MDR_COL_EXT_FORCE: return (void **)&t.ext_f;
    case MDR_COL_EXT_TORQUE: return (void **)&t.ext_t;
    case MDR_COL_LEAF_ID: return (void **)&t.leaf_id;
    default: return nullptr;
    }
}

}  // namespace mdr
