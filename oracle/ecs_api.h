/* TEST INFRASTRUCTURE ONLY -- the interface both ECS checkers export:
 *   oracle/_ref/libmadrona_ecs_ref.so  the UNMODIFIED reference ECS + task graph
 *                                      (src/core, src/common) driven by ecs_ref_harness.cpp
 *   oracle/libmadrona_ecs_port.so      the scalar restatement ecs_port.cpp
 * Worlds are independent single-world StateManagers, exactly as the reference's
 * GPU backend keeps one StateManager per world and as the physics
 * oracle does. The apps are the configs of include/madrona_b200_apps.h.
 * Nothing under gpu_ecs_madrona_b200/ links or loads either library. */
#ifndef ORACLE_ECS_API_H
#define ORACLE_ECS_API_H
#include <stdint.h>
#include "../include/madrona_b200_apps.h"

#ifdef __cplusplus
extern "C" {
#endif

const char *eorc_kind(void);
/* params: SCRIPT [with_singleton]; TASKGRAPH [num_spheres]; FANOUT [num_objs];
 * FVS [num_dragons, num_knights]. Worlds get global indices first_world.. */
void *eorc_create(int32_t app, int32_t num_worlds, int32_t first_world, const int32_t *params, int32_t nparams);
void eorc_destroy(void *h);
void eorc_step(void *h, int32_t num_steps, int32_t num_threads);
/* SCRIPT app: applies ops to `world` in order (immediate mode). */
int32_t eorc_script(void *h, int32_t world, const mdr_script_op *ops, int32_t nops);

int32_t eorc_num_archetypes(void *h);
int32_t eorc_num_cols(void *h, int32_t arch);            /* including column 0 = Entity */
int32_t eorc_col_bytes(void *h, int32_t arch, int32_t col);
int32_t eorc_num_rows(void *h, int32_t world, int32_t arch);
/* rows x bytes of one column of one world, row order as stored. */
void eorc_read_column(void *h, int32_t world, int32_t arch, int32_t col, void *out);
void eorc_write_column(void *h, int32_t world, int32_t arch, int32_t col, const void *in);
/* The entity id map of a world: per id the 12-byte node (Loc | free links,
 * generation), and state[10] = {init cache: free_head, num_free, overflow_head,
 * num_overflow; world cache: the same four; global free head; number of ids}. */
int32_t eorc_id_count(void *h, int32_t world);
void eorc_read_idmap(void *h, int32_t world, int32_t *a, int32_t *b, uint32_t *gen, int32_t state[10]);
/* Exported column `slot`: the world-major concatenation of StateManager::
 * copyOutExportedColumns (src/core/state.cpp:516-563). Returns bytes. */
int64_t eorc_export(void *h, int32_t slot, void *out, int64_t max_bytes);
/* Reference checker only: StaticIntegerMap<128>(keys, values) -> both storage arrays [128]. */
void eorc_static_map_build(const uint32_t *keys, const uint32_t *values, uint32_t n,
                           uint32_t *out_keys, uint32_t *out_values);

#ifdef __cplusplus
}
#endif
#endif
