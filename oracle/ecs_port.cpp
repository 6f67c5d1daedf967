// TEST INFRASTRUCTURE ONLY -- scalar restatement of the reference's ECS state
// layer and task graph for the configs of include/madrona_b200_apps.h, behind
// oracle/ecs_api.h. This is the checker that travels to the GPU box in source
// form. PARITY PINNED: tests/test_ecs_oracle_pin.py compares it bit for bit
// (row counts, Entity columns, every column, the whole id map including free
// links, exports) with the compiled reference (oracle/_ref/libmadrona_ecs_ref.so)
// and with tests/golden/ecs_*.npz generated from the reference.
// Never linked or loaded by gpu_ecs_madrona_b200/.
//
// Each function cites the reference code it restates (paths under /root/reference).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <thread>
#include <vector>

#include "ecs_api.h"

namespace {

constexpr int32_t kSentinel = (int32_t)0xFFFFFFFF;   // include/madrona/impl/id_map.hpp:135
constexpr int32_t kIdsPerCache = 64;                 // id_map.hpp:134

struct Entity { uint32_t gen; int32_t id; };
struct Loc { uint32_t archetype; int32_t row; };     // include/madrona/ecs.hpp:24-31
constexpr uint32_t kNoArch = 0xFFFFFFFFu;

// IDMap<Entity, Loc, ...>::Node, id_map.hpp:36-47: Loc and the free links share storage.
struct Node { int32_t a, b; uint32_t gen; };
struct Cache { int32_t free_head = kSentinel, num_free = 0, overflow_head = kSentinel, num_overflow = 0; };

struct IdMap {
    std::vector<Node> store;        // EntityStore map_(0): starts empty, src/core/state.cpp:60-62
    int32_t global_head = kSentinel;

    // id_map_impl.inl:68-112 (assignCachedID)
    Entity assign(int32_t *head)
    {
        int32_t id = *head;
        Node &n = store[id];
        int32_t run = n.b;
        if (run == 1) {
            *head = n.a;
        } else {
            Node &nx = store[id + 1];
            nx.a = n.a; nx.b = run - 1; nx.gen = 0;
            *head = id + 1;
        }
        return Entity { n.gen, id };
    }
    // id_map_impl.inl:68-186
    Entity acquire(Cache &c)
    {
        if (c.num_overflow > 0) { c.num_overflow--; return assign(&c.overflow_head); }
        if (c.num_free > 0) { c.num_free--; return assign(&c.free_head); }
        if (global_head != kSentinel) {
            int32_t h = global_head;
            global_head = store[h].b;
            store[h].b = 1;
            c.free_head = h;
            c.num_free = kIdsPerCache - 1;
            return assign(&c.free_head);
        }
        int32_t first = (int32_t)store.size();
        store.resize(store.size() + kIdsPerCache, Node { 0, 0, 0 });
        store[first].gen = 0;
        store[first + 1] = Node { kSentinel, kIdsPerCache - 1, 0 };
        c.free_head = first + 1;
        c.num_free = kIdsPerCache - 1;
        return Entity { 0, first };
    }
    // id_map_impl.inl:188-224
    void release(Cache &c, int32_t id)
    {
        Node &n = store[id];
        n.gen += 1;
        n.b = 1;
        if (c.num_free < kIdsPerCache) {
            n.a = c.free_head; c.free_head = id; c.num_free++;
            return;
        }
        if (c.num_overflow < kIdsPerCache) {
            n.a = c.overflow_head; c.overflow_head = id; c.num_overflow++;
        }
        if (c.num_overflow == kIdsPerCache) {
            store[c.overflow_head].b = global_head;
            global_head = c.overflow_head;
            c.overflow_head = kSentinel;
            c.num_overflow = 0;
        }
    }
    // id_map_impl.inl:226-331, including its quirks: the last node of every full
    // 64-block has its generation bumped twice (linkToNext at :257 then :262), and
    // linkToNext reads keys[idx + 1] one past a full block (`peek`).
    void bulkRelease(Cache &c, const Entity *keys, int64_t n, int32_t peek)
    {
        if (n <= 0) return;
        auto keyId = [&](int64_t i) { return i < n ? keys[i].id : peek; };
        auto link = [&](int64_t i) {
            Node &nd = store[keys[i].id];
            nd.gen += 1; nd.a = keyId(i + 1); nd.b = 1;
        };
        int64_t base, rem = 0;
        int32_t tail = kSentinel;   // id of global_tail_node
        for (base = 0; base < n; base += kIdsPerCache) {
            rem = n - base;
            if (rem < kIdsPerCache) break;
            int32_t head_id = keys[base].id;
            for (int s = 0; s < kIdsPerCache; s++) link(base + s);
            Node &last = store[keys[base + kIdsPerCache - 1].id];
            last.gen += 1; last.a = kSentinel; last.b = 1;
            if (tail != kSentinel) store[tail].b = head_id;
            tail = head_id;
        }
        if (rem != kIdsPerCache) {
            int32_t start = keys[base].id;
            for (int64_t i = base; i < n - 1; i++) link(i);
            Node &t = store[keys[n - 1].id];
            t.gen += 1; t.b = 1; t.a = c.overflow_head;
            int64_t from_overflow = kIdsPerCache - rem;
            if (c.num_overflow < from_overflow) {
                c.overflow_head = start;
                c.num_overflow += (int32_t)rem;
            } else {
                int32_t next = c.overflow_head, cut = kSentinel;
                for (int64_t i = 0; i < from_overflow; i++) { cut = next; next = store[cut].a; }
                store[cut].a = kSentinel;
                c.overflow_head = next;
                c.num_overflow -= (int32_t)from_overflow;
                if (tail != kSentinel) store[tail].b = start;
                tail = start;
            }
        }
        if (tail == kSentinel) return;
        store[tail].b = global_head;
        global_head = keys[0].id;
    }
    Loc lookup(Entity e) const   // id_map.hpp:59-68
    {
        if (e.id < 0 || e.id >= (int32_t)store.size() || store[e.id].gen != e.gen) return Loc { kNoArch, 0 };
        return Loc { (uint32_t)store[e.id].a, store[e.id].b };
    }
};

// Table, src/common/table.cpp:19-86: one byte vector per column. Growth policy
// (x2, min 10) does not affect observable state and is left to std::vector,
// except that rows past num_rows keep their old bytes as realloc does.
struct Table {
    std::vector<int32_t> col_bytes;
    std::vector<std::vector<uint8_t>> cols;
    int32_t num_rows = 0;
    uint8_t *at(int c, int32_t r) { return cols[c].data() + (size_t)r * col_bytes[c]; }
    int32_t addRow()
    {
        int32_t r = num_rows++;
        for (size_t c = 0; c < cols.size(); c++)
            if (cols[c].size() < (size_t)num_rows * col_bytes[c]) cols[c].resize((size_t)num_rows * col_bytes[c] * 2 + 64);
        return r;
    }
    bool removeRow(int32_t row)   // table.cpp:59-71
    {
        int32_t from = --num_rows;
        if (from == row) return false;
        for (size_t c = 0; c < cols.size(); c++) memcpy(at((int)c, row), at((int)c, from), col_bytes[c]);
        return true;
    }
};

struct World {
    uint32_t index = 0;
    IdMap ids;
    Cache init_cache, cache;     // StateManager::init_state_cache_ vs the worker's StateCache
    std::vector<Table> tables;   // archetype id -> table; column 0 = Entity (state.cpp:401-404, non-MW)
    std::vector<Entity> handles;
    int32_t peekId(Table &t, int64_t n)   // the Entity bytes one past the last row, as the reference reads them
    {
        if ((size_t)(n + 1) * 8 > t.cols[0].size()) return 0;
        Entity e; memcpy(&e, t.at(0, (int32_t)n), 8);
        return e.id;
    }
    // include/madrona/state.inl:398-449
    Entity makeEntityNow(uint32_t arch, Cache &c)
    {
        Entity e = ids.acquire(c);
        Table &t = tables[arch];
        int32_t row = t.addRow();
        memcpy(t.at(0, row), &e, 8);
        ids.store[e.id].a = (int32_t)arch;
        ids.store[e.id].b = row;
        return e;
    }
    Loc makeTemporary(uint32_t arch) { return Loc { arch, tables[arch].addRow() }; }   // state.inl:451-464
    // src/core/state.cpp:181-202
    void destroyEntityNow(Entity e)
    {
        Loc loc = ids.lookup(e);
        if (loc.archetype == kNoArch) return;
        Table &t = tables[loc.archetype];
        if (t.removeRow(loc.row)) {
            Entity moved; memcpy(&moved, t.at(0, loc.row), 8);
            ids.store[moved.id].b = loc.row;
        }
        ids.release(cache, e.id);
    }
    // src/core/state.cpp:565-581
    void clear(uint32_t arch, bool is_temporary)
    {
        Table &t = tables[arch];
        if (!is_temporary && t.num_rows > 0)
            ids.bulkRelease(cache, (const Entity *)t.cols[0].data(), t.num_rows, peekId(t, t.num_rows));
        t.num_rows = 0;
    }
    template <typename T> T *col(uint32_t arch, int c) { return (T *)tables[arch].cols[c].data(); }
};

struct Export { uint32_t arch, col; };
struct Handle {
    int app = 0;
    std::vector<World> worlds;
    std::vector<Export> exports;
    std::vector<int32_t> params;
};

void addArchetype(World &w, std::initializer_list<int32_t> user_bytes)
{
    Table t;
    t.col_bytes.push_back(8);
    for (int32_t b : user_bytes) t.col_bytes.push_back(b);
    t.cols.resize(t.col_bytes.size());
    w.tables.push_back(std::move(t));
}
// registerSingleton, include/madrona/state.inl:171-187: a one-row archetype whose
// entity is made through the StateManager's own init cache.
void addSingleton(World &w, int32_t bytes)
{
    addArchetype(w, { bytes });
    w.makeEntityNow((uint32_t)w.tables.size() - 1, w.init_cache);
}

// ------------------------------------------------------------------ systems --
// Iteration order of a query = archetype id order, then row order
// (src/core/state.cpp:270-361, include/madrona/state.inl:358-396); node order =
// insertion order of a dependency chain (src/core/taskgraph.cpp:39-109).
void stepTaskgraph(World &w)
{
    for (uint32_t a = 0; a < 2; a++) {
        mdr_vec3 *p = w.col<mdr_vec3>(a, 1);
        for (int32_t i = 0; i < w.tables[a].num_rows; i++) mdr_tg_drift(&p[i]);
    }
    for (uint32_t a = 0; a < 2; a++) {
        mdr_vec3 *p = w.col<mdr_vec3>(a, 1);
        for (int32_t i = 0; i < w.tables[a].num_rows; i++) mdr_tg_clamp(&p[i]);
    }
}

void stepFanout(World &w)
{
    w.clear(1, true);
    Table &t = w.tables[0];
    const int32_t n = t.num_rows;
    for (int32_t i = 0; i < n; i++) mdr_fan_move(&w.col<mdr_vec3>(0, 1)[i]);
    for (int32_t i = 0; i < n; i++)
        mdr_fan_box(&w.col<mdr_vec3>(0, 1)[i], &w.col<mdr_quat>(0, 2)[i], &w.col<mdr_box>(0, 3)[i]);
    for (int32_t i = 0; i < n; i++) {
        Entity e = w.col<Entity>(0, 0)[i];
        const int32_t row = w.ids.lookup(e).row;
        for (int32_t j = row + 1; j < n; j++) {
            if (!mdr_box_overlap(&w.col<mdr_box>(0, 3)[i], &w.col<mdr_box>(0, 3)[j])) continue;
            Loc l = w.makeTemporary(1);
            Entity eb = w.col<Entity>(0, 0)[j];
            mdr_candidate c { { e.gen, e.id }, { eb.gen, eb.id } };
            w.col<mdr_candidate>(1, 1)[l.row] = c;
            w.col<uint32_t>(1, 2)[l.row] = mdr_fan_key(row, j);
        }
    }
    for (int32_t k = 0; k < w.tables[1].num_rows; k++) {
        mdr_candidate c = w.col<mdr_candidate>(1, 1)[k];
        Loc la = w.ids.lookup(Entity { c.a.gen, c.a.id }), lb = w.ids.lookup(Entity { c.b.gen, c.b.id });
        w.col<int32_t>(la.archetype, 4)[la.row] += 1;
        w.col<int32_t>(lb.archetype, 4)[lb.row] += 1;
    }
}

void stepFvs(World &w)
{
    const uint32_t tick = w.col<uint32_t>(3, 1)[0];
    for (uint32_t a = 0; a < 2; a++)
        for (int32_t i = 0; i < w.tables[a].num_rows; i++)
            mdr_fvs_action_select(w.index, tick, w.col<Entity>(a, 0)[i].id, &w.col<mdr_vec3>(a, 1)[i],
                                  &w.col<float>(a, 3)[i]);
    for (int32_t i = 0; i < w.tables[0].num_rows; i++) {     // Query<Entity, Action, Mana>: Dragon only
        mdr_vec3 target;
        if (!mdr_fvs_caster(w.index, tick, w.col<Entity>(0, 0)[i].id, &w.col<float>(0, 3)[i],
                            &w.col<float>(0, 4)[i], &target))
            continue;
        for (uint32_t a = 0; a < 2; a++)
            for (int32_t k = 0; k < w.tables[a].num_rows; k++)
                if (mdr_fvs_in_blast(&target, &w.col<mdr_vec3>(a, 1)[k])) w.col<int32_t>(a, 2)[k] -= MDR_FVS_BLAST_DAMAGE;
    }
    for (int32_t i = 0; i < w.tables[1].num_rows; i++) {     // Query<Entity, Action, Quiver>: Knight only
        int32_t target = mdr_fvs_archer(w.index, tick, w.col<Entity>(1, 0)[i].id, &w.col<float>(1, 3)[i],
                                        &w.col<int32_t>(1, 4)[i], w.tables[0].num_rows);
        if (target >= 0) w.col<int32_t>(0, 2)[target] -= MDR_FVS_ARROW_DAMAGE;
    }
    for (uint32_t a = 0; a < 2; a++)
        for (int32_t i = 0; i < w.tables[a].num_rows; i++) {
            if (w.col<int32_t>(a, 2)[i] > 0) continue;
            Entity e = w.col<Entity>(a, 0)[i];
            Loc l = w.makeTemporary(2);
            w.col<mdr_cleanup>(2, 1)[l.row] = mdr_cleanup { { e.gen, e.id }, (int32_t)a };
            w.col<uint32_t>(2, 2)[l.row] = mdr_fvs_cleanup_key(a, i);
        }
    for (int32_t k = 0; k < w.tables[2].num_rows; k++) {
        mdr_cleanup c = w.col<mdr_cleanup>(2, 1)[k];
        w.destroyEntityNow(Entity { c.e.gen, c.e.id });
    }
    for (int32_t k = 0; k < w.tables[2].num_rows; k++) {
        mdr_cleanup c = w.col<mdr_cleanup>(2, 1)[k];
        mdr_vec3 p; int32_t hp, arrows; float mp;
        mdr_fvs_spawn(w.index, tick, c.e.id, c.kind, &p, &hp, &mp, &arrows);
        Entity e = w.makeEntityNow((uint32_t)c.kind, w.cache);
        Loc l = w.ids.lookup(e);
        w.col<mdr_vec3>(l.archetype, 1)[l.row] = p;
        w.col<int32_t>(l.archetype, 2)[l.row] = hp;
        w.col<float>(l.archetype, 3)[l.row] = 0.f;
        if (c.kind == MDR_FVS_DRAGON) w.col<float>(0, 4)[l.row] = mp;
        else w.col<int32_t>(1, 4)[l.row] = arrows;
    }
    w.clear(2, true);
    w.col<uint32_t>(3, 1)[0] = tick + 1;
}

void stepWorld(Handle *h, World &w)
{
    switch (h->app) {
    case MDR_APP_TASKGRAPH: stepTaskgraph(w); break;
    case MDR_APP_FANOUT: stepFanout(w); break;
    case MDR_APP_FVS: stepFvs(w); break;
    default: break;
    }
}

}  // namespace

extern "C" {

const char *eorc_kind(void) { return "port"; }

void *eorc_create(int32_t app, int32_t num_worlds, int32_t first_world, const int32_t *params, int32_t nparams)
{
    if (app < 0 || app >= MDR_APP_CHURN || num_worlds <= 0) return nullptr;   // churn: compiled reference only
    Handle *h = new Handle;
    h->app = app;
    h->params.assign(params, params + (params ? nparams : 0));
    auto P = [&](int i, int32_t dflt) { return i < nparams ? params[i] : dflt; };
    h->worlds.resize(num_worlds);
    for (int wi = 0; wi < num_worlds; wi++) {
        World &w = h->worlds[wi];
        w.index = (uint32_t)(first_world + wi);
        switch (app) {
        case MDR_APP_SCRIPT:
            addArchetype(w, { 4 }); addArchetype(w, { 4, 12 }); addArchetype(w, { 4, 12, 16 }); addArchetype(w, { 4 });
            if (P(0, 1)) { addSingleton(w, 4); memset(w.tables[4].at(1, 0), 0, 4); }
            break;
        case MDR_APP_TASKGRAPH:
            addArchetype(w, { 12, 16 }); addArchetype(w, { 12, 16 });
            for (int i = 0, n = P(0, 100); i < n; i++) w.makeEntityNow(0, w.cache);
            w.makeEntityNow(1, w.cache);
            h->exports = { { 1, 1 }, { 1, 2 } };
            break;
        case MDR_APP_FANOUT:
            addArchetype(w, { 12, 16, 24, 4 }); addArchetype(w, { 16, 4 });
            for (int i = 0, n = P(0, 100); i < n; i++) w.makeEntityNow(0, w.cache);
            h->exports = { { 0, 4 } };
            break;
        case MDR_APP_FVS:
            addArchetype(w, { 12, 4, 4, 4 }); addArchetype(w, { 12, 4, 4, 4 }); addArchetype(w, { 12, 4 });
            addSingleton(w, 4);
            memset(w.tables[3].at(1, 0), 0, 4);
            for (int i = 0, n = P(0, 50); i < n; i++) w.makeEntityNow(0, w.cache);
            for (int i = 0, n = P(1, 200); i < n; i++) w.makeEntityNow(1, w.cache);
            h->exports = { { 0, 2 }, { 1, 1 } };
            break;
        }
    }
    return h;
}

void eorc_destroy(void *hp) { delete (Handle *)hp; }

void eorc_step(void *hp, int32_t num_steps, int32_t num_threads)
{
    Handle *h = (Handle *)hp;
    const int W = (int)h->worlds.size();
    if (num_threads <= 1) {
        for (int s = 0; s < num_steps; s++)
            for (int i = 0; i < W; i++) stepWorld(h, h->worlds[i]);
        return;
    }
    std::atomic<int> next { 0 };   // src/mw/cpu_exec.cpp:244-284
    std::vector<std::thread> pool;
    for (int t = 0; t < num_threads; t++)
        pool.emplace_back([&]() {
            while (true) {
                int i = next.fetch_add(1, std::memory_order_relaxed);
                if (i >= W) break;
                for (int s = 0; s < num_steps; s++) stepWorld(h, h->worlds[i]);
            }
        });
    for (auto &t : pool) t.join();
}

int32_t eorc_script(void *hp, int32_t world, const mdr_script_op *ops, int32_t nops)
{
    Handle *h = (Handle *)hp;
    if (h->app != MDR_APP_SCRIPT) return -1;
    World &w = h->worlds[world];
    for (int i = 0; i < nops; i++) {
        const mdr_script_op &o = ops[i];
        switch (o.op) {
        case MDR_OP_CREATE: {
            if (o.a < 0 || o.a > 2) return -1;
            float b[3], c[4];
            mdr_script_payload(o.b, b, c);
            Entity e = w.makeEntityNow((uint32_t)o.a, w.cache);
            Loc l = w.ids.lookup(e);
            w.col<uint32_t>(l.archetype, 1)[l.row] = (uint32_t)o.b;
            if (o.a >= 1) memcpy(w.tables[o.a].at(2, l.row), b, 12);
            if (o.a >= 2) memcpy(w.tables[o.a].at(3, l.row), c, 16);
            w.handles.push_back(e);
        } break;
        case MDR_OP_DESTROY:
            if (o.a >= 0 && o.a < (int)w.handles.size()) w.destroyEntityNow(w.handles[o.a]);
            break;
        case MDR_OP_CLEAR:
            if (o.a < 0 || o.a > 2) return -1;
            w.clear((uint32_t)o.a, false);
            break;
        case MDR_OP_TMP: {
            Loc l = w.makeTemporary(3);
            w.col<uint32_t>(3, 1)[l.row] = (uint32_t)o.b;
        } break;
        case MDR_OP_CLEAR_TMP:
            w.clear(3, true);
            break;
        default:
            return -1;
        }
    }
    return 0;
}

int32_t eorc_num_archetypes(void *hp) { return (int32_t)((Handle *)hp)->worlds[0].tables.size(); }
int32_t eorc_num_cols(void *hp, int32_t arch) { return (int32_t)((Handle *)hp)->worlds[0].tables[arch].col_bytes.size(); }
int32_t eorc_col_bytes(void *hp, int32_t arch, int32_t col) { return ((Handle *)hp)->worlds[0].tables[arch].col_bytes[col]; }
int32_t eorc_num_rows(void *hp, int32_t world, int32_t arch) { return ((Handle *)hp)->worlds[world].tables[arch].num_rows; }

void eorc_read_column(void *hp, int32_t world, int32_t arch, int32_t col, void *out)
{
    Table &t = ((Handle *)hp)->worlds[world].tables[arch];
    memcpy(out, t.cols[col].data(), (size_t)t.num_rows * t.col_bytes[col]);
}

void eorc_write_column(void *hp, int32_t world, int32_t arch, int32_t col, const void *in)
{
    Table &t = ((Handle *)hp)->worlds[world].tables[arch];
    memcpy(t.cols[col].data(), in, (size_t)t.num_rows * t.col_bytes[col]);
}

int32_t eorc_id_count(void *hp, int32_t world) { return (int32_t)((Handle *)hp)->worlds[world].ids.store.size(); }

void eorc_read_idmap(void *hp, int32_t world, int32_t *a, int32_t *b, uint32_t *gen, int32_t state[10])
{
    World &w = ((Handle *)hp)->worlds[world];
    for (size_t i = 0; i < w.ids.store.size(); i++) {
        a[i] = w.ids.store[i].a; b[i] = w.ids.store[i].b; gen[i] = w.ids.store[i].gen;
    }
    const Cache *cs[2] = { &w.init_cache, &w.cache };
    for (int k = 0; k < 2; k++) {
        state[4 * k] = cs[k]->free_head; state[4 * k + 1] = cs[k]->num_free;
        state[4 * k + 2] = cs[k]->overflow_head; state[4 * k + 3] = cs[k]->num_overflow;
    }
    state[8] = w.ids.global_head;
    state[9] = (int32_t)w.ids.store.size();
}

int64_t eorc_export(void *hp, int32_t slot, void *out, int64_t max_bytes)
{
    Handle *h = (Handle *)hp;
    if (slot < 0 || slot >= (int)h->exports.size()) return -1;
    int64_t off = 0;
    for (World &w : h->worlds) {   // src/core/state.cpp:516-563
        Table &t = w.tables[h->exports[slot].arch];
        int64_t nb = (int64_t)t.num_rows * t.col_bytes[h->exports[slot].col];
        if (off + nb > max_bytes) return -1;
        memcpy((char *)out + off, t.cols[h->exports[slot].col].data(), nb);
        off += nb;
    }
    return off;
}

}  // extern "C"
