// Widened-id component storage shared with the device.
//
//   reference                                                  here (namespace icx_host)
//   pleep::ComponentArray<T>    ecs/component_array.h:62-70    ComponentPools (the physics signature: Transform, Physics,
//     packed array + Entity <-> index maps, 16-bit Entity        ColliderComponent; one row per entity, 32-bit widened Entity)
//   pleep::I_Synchro            ecs/i_synchro.h:48             rows are kept in ascending id order (std::set<Entity>)
//   PhysicsSynchro / ColliderSynchro::update                   PhysicsDynamo::bind(pools): no per-entity packets, no copies of
//     physics_synchro.cpp:29-35, collider_synchro.cpp:28-42      unchanged rows, see below
//
// The reference's dynamo works on references into the component arrays, so "submitting" costs nothing there.  A GPU tick
// needs the rows on the device; what keeps that cheap is that the device copy is authoritative between ticks and the host
// says what it touched:
//   * rows live in page-locked memory (icx_host_alloc): every transfer is one asynchronous DMA at PCIe rate;
//   * edit_transform / edit_physics / edit_colliders hand out a mutable row AND mark it: the next tick sends only marked
//     rows (icx_bodies_update with a row list), or everything when more than a quarter is marked;
//   * insert / remove keep the ascending order; the device follows through icx_entities_create / icx_entities_remove, so the
//     rows that stay never cross PCIe again;
//   * after the tick the Transform + Physics rows come back in place (icx_bodies_download straight into the pinned rows).
// Header-only; link with libicx.so.
#pragma once
#include <algorithm>
#include <cstring>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/icx.h"

namespace icx_host {

class ComponentPools {
public:
    explicit ComponentPools(size_t capacity) : m_cap(capacity)
    {
        m_ids = alloc<IcxEntity>(); m_T = alloc<IcxTransform>(); m_P = alloc<IcxPhysics>(); m_C = alloc<IcxColliderComponent>();
        m_hasP = alloc<uint8_t>(); m_ts = alloc<uint8_t>(); m_tc = alloc<uint16_t>();
        m_dirty.assign(capacity, 0);
    }
    ~ComponentPools() { for (void* p : {(void*)m_ids, (void*)m_T, (void*)m_P, (void*)m_C, (void*)m_hasP, (void*)m_ts, (void*)m_tc}) icx_host_free(p); }
    ComponentPools(const ComponentPools&) = delete;
    ComponentPools& operator=(const ComponentPools&) = delete;

    size_t size() const { return m_n; }
    size_t capacity() const { return m_cap; }
    static constexpr size_t npos = (size_t)-1;
    // ComponentArray::m_mapEntityToIndex, as a binary search over the ascending ids
    size_t row_of(IcxEntity id) const
    {
        const IcxEntity* e = std::lower_bound(m_ids, m_ids + m_n, id);
        return (e != m_ids + m_n && *e == id) ? (size_t)(e - m_ids) : npos;
    }
    IcxEntity entity(size_t row) const { return m_ids[row]; }

    // Cosmos::register_entity + add_component for the physics signature.  Returns the row; an id that exists is left alone.
    size_t insert(IcxEntity id, const IcxTransform& t, const IcxPhysics* p, const IcxColliderComponent& c, uint8_t tstate = ICX_TS_MERGED, uint16_t tcoh = 0)
    {
        size_t at = (size_t)(std::lower_bound(m_ids, m_ids + m_n, id) - m_ids);
        if (at < m_n && m_ids[at] == id) return at;
        if (m_n == m_cap) throw std::length_error("ComponentPools: capacity exhausted");
        shift(at, m_n, +1);
        m_ids[at] = id; m_T[at] = t; m_hasP[at] = p ? 1 : 0;
        if (p) m_P[at] = *p; else std::memset(&m_P[at], 0, sizeof(IcxPhysics));
        m_C[at] = c; m_ts[at] = tstate; m_tc[at] = tcoh;
        m_dirty.insert(m_dirty.begin() + at, 0); m_dirty.pop_back();
        m_n++;
        if (m_synced) m_created.push_back(id); else m_structure = true;
        return at;
    }
    // Cosmos::destroy_entity.  false if the id is not there.
    bool remove(IcxEntity id)
    {
        const size_t at = row_of(id);
        if (at == npos) return false;
        shift(at + 1, m_n, -1);
        m_dirty.erase(m_dirty.begin() + at); m_dirty.push_back(0);
        m_n--;
        auto pending = std::find(m_created.begin(), m_created.end(), id);
        if (pending != m_created.end()) m_created.erase(pending);                   // never reached the device
        else if (m_synced) m_removed.push_back(id); else m_structure = true;
        return true;
    }

    // read access (ComponentArray::get_data_for) ...
    const IcxTransform& transform(size_t row) const { return m_T[row]; }
    const IcxPhysics& physics(size_t row) const { return m_P[row]; }
    const IcxColliderComponent& colliders(size_t row) const { return m_C[row]; }
    bool has_physics(size_t row) const { return m_hasP[row] != 0; }
    uint8_t timestream_state(size_t row) const { return m_ts[row]; }
    uint16_t timestream_coherency(size_t row) const { return m_tc[row]; }
    // ... and write access, which marks the row for the next tick
    IcxTransform& edit_transform(size_t row) { m_dirty[row] |= 1; return m_T[row]; }
    IcxPhysics& edit_physics(size_t row) { m_dirty[row] |= 2; return m_P[row]; }
    // changing which collider slots are active changes the packet structure: the whole pool is resubmitted
    IcxColliderComponent& edit_colliders(size_t row) { m_dirty[row] |= 4; return m_C[row]; }
    void set_timestream_state(size_t row, uint8_t state, uint16_t coherency) { m_ts[row] = state; m_tc[row] = coherency; m_structure = true; }
    void mark_all() { std::fill(m_dirty.begin(), m_dirty.begin() + m_n, 3); }

    struct SyncStats { size_t rows_sent = 0, bytes_h2d = 0, bytes_d2h = 0; bool full_upload = false; };
    const SyncStats& last_sync() const { return m_stats; }

    // Host -> device before a tick: what changed since the last call, nothing else.
    void to_device(IcxWorld* w)
    {
        m_stats = SyncStats{};
        const uint32_t n = (uint32_t)m_n;
        if (!m_synced || m_structure) { full_upload(w); return; }
        if (!m_removed.empty()) {
            uint32_t k = 0;
            ck(icx_entities_remove(w, (uint32_t)m_removed.size(), m_removed.data(), &k), w, "icx_entities_remove");
            m_removed.clear();
        }
        if (!m_created.empty()) {                                                   // the new rows only; the rest stays on the device
            const size_t k = m_created.size();
            std::vector<IcxTransform> t(k); std::vector<IcxPhysics> p(k); std::vector<IcxColliderComponent> c(k);
            std::vector<uint8_t> hp(k), ts(k); std::vector<uint16_t> tc(k);
            for (size_t i = 0; i < k; i++) {
                const size_t r = row_of(m_created[i]);
                t[i] = m_T[r]; p[i] = m_P[r]; c[i] = m_C[r]; hp[i] = m_hasP[r]; ts[i] = m_ts[r]; tc[i] = m_tc[r]; m_dirty[r] = 0;
            }
            uint32_t made = 0;
            ck(icx_entities_create(w, (uint32_t)k, m_created.data(), t.data(), p.data(), hp.data(), c.data(), ts.data(), tc.data(), &made), w, "icx_entities_create");
            m_stats.bytes_h2d += k * (sizeof(IcxTransform) + sizeof(IcxPhysics) + sizeof(IcxColliderComponent) + 8);
            m_created.clear();
        }
        m_rows.clear();
        uint8_t any = 0;
        for (uint32_t r = 0; r < n; r++) if (m_dirty[r]) { m_rows.push_back(r); any |= m_dirty[r]; }
        if (m_rows.empty()) return;
        const size_t k = m_rows.size();
        int rc;
        if (k * 4 > (size_t)n && !(any & 4)) {                                      // most rows touched: one DMA per array beats a gather
            rc = icx_bodies_update(w, n, nullptr, (any & 1) ? m_T : nullptr, (any & 2) ? m_P : nullptr, nullptr);
            m_stats.rows_sent = n; m_stats.bytes_h2d += (size_t)n * (((any & 1) ? sizeof(IcxTransform) : 0) + ((any & 2) ? sizeof(IcxPhysics) : 0));
        } else {
            m_sT.resize((any & 1) ? k : 0); m_sP.resize((any & 2) ? k : 0); m_sC.resize((any & 4) ? k : 0);
            for (size_t i = 0; i < k; i++) {
                const uint32_t r = m_rows[i];
                if (any & 1) m_sT[i] = m_T[r];
                if (any & 2) m_sP[i] = m_P[r];
                if (any & 4) m_sC[i] = m_C[r];
            }
            rc = icx_bodies_update(w, (uint32_t)k, m_rows.data(), (any & 1) ? m_sT.data() : nullptr, (any & 2) ? m_sP.data() : nullptr, (any & 4) ? m_sC.data() : nullptr);
            m_stats.rows_sent = k;
            m_stats.bytes_h2d += k * (4 + ((any & 1) ? sizeof(IcxTransform) : 0) + ((any & 2) ? sizeof(IcxPhysics) : 0) + ((any & 4) ? sizeof(IcxColliderComponent) : 0));
        }
        if (rc == ICX_E_STATE) { full_upload(w); return; }                          // a collider slot was switched on / off: new packet structure
        ck(rc, w, "icx_bodies_update");
        for (uint32_t r : m_rows) m_dirty[r] = 0;
    }

    // Device -> host after a tick: Transform + Physics rows (and the timestream states the tick may have changed) in place.
    void from_device(IcxWorld* w, bool colliders = false)
    {
        const uint32_t n = (uint32_t)m_n;
        ck(icx_bodies_download(w, n, m_T, m_P, colliders ? m_C : nullptr, m_ts, m_tc), w, "icx_bodies_download");
        m_stats.bytes_d2h = (size_t)n * (sizeof(IcxTransform) + sizeof(IcxPhysics) + 3 + (colliders ? sizeof(IcxColliderComponent) : 0));
    }

private:
    template <typename T> T* alloc()
    {
        T* p = (T*)icx_host_alloc(std::max<size_t>(m_cap, 1) * sizeof(T));
        if (!p) throw std::bad_alloc();                                             // no CUDA device: there is no CPU fallback
        std::memset(p, 0, std::max<size_t>(m_cap, 1) * sizeof(T));
        return p;
    }
    template <typename T> static void move_rows(T* a, size_t from, size_t to, long by) { std::memmove(a + from + by, a + from, (to - from) * sizeof(T)); }
    void shift(size_t from, size_t to, long by)
    {
        if (from >= to) return;
        move_rows(m_ids, from, to, by); move_rows(m_T, from, to, by); move_rows(m_P, from, to, by); move_rows(m_C, from, to, by);
        move_rows(m_hasP, from, to, by); move_rows(m_ts, from, to, by); move_rows(m_tc, from, to, by);
    }
    void full_upload(IcxWorld* w)
    {
        const uint32_t n = (uint32_t)m_n;
        ck(icx_bodies_upload(w, n, m_ids, m_T, m_P, m_hasP, m_C, m_ts, m_tc), w, "icx_bodies_upload");
        std::fill(m_dirty.begin(), m_dirty.end(), 0);
        m_created.clear(); m_removed.clear();
        m_synced = true; m_structure = false;
        m_stats.full_upload = true; m_stats.rows_sent = n;
        m_stats.bytes_h2d = (size_t)n * (sizeof(IcxEntity) + sizeof(IcxTransform) + sizeof(IcxPhysics) + sizeof(IcxColliderComponent) + 4);
    }
    static void ck(int rc, IcxWorld* w, const char* what)
    {
        if (rc != ICX_OK) throw std::runtime_error(std::string(what) + ": " + icx_status_string(rc) + " -- " + (w ? icx_last_error(w) : ""));
    }

    size_t m_cap, m_n = 0;
    IcxEntity* m_ids; IcxTransform* m_T; IcxPhysics* m_P; IcxColliderComponent* m_C; uint8_t* m_hasP; uint8_t* m_ts; uint16_t* m_tc;
    std::vector<uint8_t> m_dirty;                       // per row: 1 transform, 2 physics, 4 colliders edited since the last to_device
    std::vector<IcxEntity> m_created, m_removed;        // entity lifecycle since the last to_device
    std::vector<uint32_t> m_rows; std::vector<IcxTransform> m_sT; std::vector<IcxPhysics> m_sP; std::vector<IcxColliderComponent> m_sC;
    bool m_synced = false, m_structure = false;
    SyncStats m_stats;
};

} // namespace icx_host
