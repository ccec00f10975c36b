// Host-side mirror of the timestream half of pleep::TimelineApi (source/networking/timeline_api.h:53-107)
// for the dense ENTITY_UPDATE stream, forwarding to the device snapshot ring of the C ABI.
//
//   reference (per entity, per tick, heap message + 2 mutexes)      here (one call per tick per cosmos)
//   push_past_timestream(entity, msg)       timeline_api.cpp:98-108   push_past_timestream(world, coherency)
//   pop_future_timestream(entity, coh, msg) timeline_api.cpp:110-121  pop_future_timestream(world, coherency)
//   push_timestream_at_breakpoint(...)      timeline_api.cpp:123-133  push_timestream_at_breakpoint(world, coherency)
//   pop_timestream_at_breakpoint(...)       timeline_api.cpp:135-146  pop_timestream_at_breakpoint(world, coherency)
//   link_timestreams (sets breakpoints)     timeline_api.cpp:148-172  link_timestreams() / unlink_timestreams()
//   the ENTITY_CREATED / REMOVED, JUMP_* messages of the same       push_past_event(coherency, records) /
//   streams (server_network_dynamo.cpp:748-811, :878-976, :422-504)   pop_future_events(now) -> records; recorded_entity(...)
// Event records ride in the tick's snapshot slot (and in its transport buffer); their handlers stay host code and call
// Cosmos::register_entity / condemn_entity = create_entities / remove_entities below.
#pragma once
#include <vector>
#include "physics_dynamo.hpp"

namespace icx_host {

class TimelineApi {
public:
    // delay_ticks = timesliceDelay * FRAMERATE (timeline_config.h:34, i_cosmos_context.h:24): 5 s * 45 Hz = 225
    TimelineApi(uint32_t max_entities, uint32_t delay_ticks = 225, int device = 0)
    {
        const int rc = icx_timeline_create(device, delay_ticks + 31, max_entities, &m_tl);
        if (rc != ICX_OK) throw std::runtime_error(std::string("icx_timeline_create: ") + icx_status_string(rc));
    }
    ~TimelineApi() { icx_timeline_destroy(m_tl); }
    TimelineApi(const TimelineApi&) = delete;
    TimelineApi& operator=(const TimelineApi&) = delete;

    void push_past_timestream(PhysicsDynamo& cosmos, uint16_t coherency) { ck(icx_record(cosmos.world(), m_tl, coherency)); }
    // true if a record stamped `coherency` was applied to the non-divergent entities
    bool pop_future_timestream(PhysicsDynamo& cosmos, uint16_t coherency)
    {
        int applied = 0;
        ck(icx_replay(cosmos.world(), m_tl, coherency, ICX_REPLAY_SERVER, &applied));
        return applied != 0;
    }
    void push_timestream_at_breakpoint(PhysicsDynamo& cosmos, uint16_t coherency) { ck(icx_record_at_breakpoint(cosmos.world(), m_tl, coherency)); }
    void pop_timestream_at_breakpoint(PhysicsDynamo& cosmos, uint16_t coherency)
    {
        ck(icx_replay(cosmos.world(), m_tl, coherency, ICX_REPLAY_PARALLEL, nullptr));
    }
    // ---- event records (ENTITY_CREATED / REMOVED, JUMP_REQUEST / DEPARTURE / ARRIVAL) ----
    void push_past_event(uint16_t coherency, const IcxTimelineEvent* records, uint32_t count) { ck(icx_timeline_event_push(m_tl, coherency, records, count)); }
    // what has come due at `now`, in stream order, stale records dropped; call it before pop_future_timestream of the same tick
    std::vector<IcxTimelineEvent> pop_future_events(uint16_t now, bool at_breakpoint = false)
    {
        std::vector<IcxTimelineEvent> out(4 * ICX_TIMELINE_EVENTS_PER_TICK);
        size_t n = 0;
        ck(icx_timeline_events_pop(m_tl, now, at_breakpoint ? ICX_REPLAY_PARALLEL : ICX_REPLAY_SERVER, out.data(), out.size(), &n));
        out.resize(std::min(n, out.size()));
        return out;
    }
    // the components the tick's record holds for `entity` (what a JUMP_ARRIVAL message carries in the reference, :940-976);
    // false if the record does not hold it.  *has_collider: the record carried its ColliderComponent too.
    bool recorded_entity(uint16_t coherency, Entity entity, TransformComponent& t, PhysicsComponent& p, bool& has_physics, ColliderComponent& c, bool* has_collider = nullptr)
    {
        int found = 0; uint8_t hp = 0;
        ck(icx_timeline_entity_fetch(m_tl, coherency, entity, &t, &p, &hp, &c, &found));
        has_physics = hp != 0;
        if (has_collider) *has_collider = (found & 2) != 0;
        return (found & 1) != 0;
    }
    // Cosmos::register_entity + components / Cosmos::condemn_entity on the device world: the entities it holds stay there
    static uint32_t create_entities(PhysicsDynamo& cosmos, uint32_t count, const Entity* ids, const TransformComponent* t, const PhysicsComponent* p,
                                    const uint8_t* has_physics, const ColliderComponent* c, const uint8_t* tstate = nullptr, const uint16_t* tcoh = nullptr)
    {
        uint32_t made = 0;
        check(icx_entities_create(cosmos.world(), count, ids, t, p, has_physics, c, tstate, tcoh, &made), cosmos.world(), "icx_entities_create");
        return made;
    }
    static uint32_t remove_entities(PhysicsDynamo& cosmos, uint32_t count, const Entity* ids)
    {
        uint32_t gone = 0;
        check(icx_entities_remove(cosmos.world(), count, ids, &gone), cosmos.world(), "icx_entities_remove");
        return gone;
    }

    void link_timestreams() { ck(icx_breakpoint_set(m_tl)); }
    void unlink_timestreams() { ck(icx_breakpoint_clear(m_tl)); }

    // fork ageing of ServerNetworkDynamo::run_relays phase 5 (server_network_dynamo.cpp:681-705);
    // returns true when parallel_notify_divergence() is due
    static bool age_forks(PhysicsDynamo& cosmos)
    {
        int need = 0;
        check(icx_fork_age(cosmos.world(), nullptr, &need), cosmos.world(), "icx_fork_age");
        return need != 0;
    }
    // congruence of the live cosmos against the record stamped `coherency` (TimejumpConditions::operator==)
    bool is_congruent(PhysicsDynamo& cosmos, uint16_t coherency, uint32_t* incongruent = nullptr, double* score = nullptr)
    {
        uint8_t flag = 0;
        ck(icx_divergence_scan(cosmos.world(), m_tl, coherency, 1, &flag, incongruent, score));
        return flag == 0;
    }
    // ParallelCosmosContext::load_and_link (parallel_cosmos_context.cpp:60-169): `parallel` becomes the copy of
    // `source` the parallel cosmos simulates (entities with chain-link 0 and no future timestream are left out,
    // forked entities carry their state and are re-merged in the source); returns the number of entities copied.
    // The caller links the timestreams afterwards (link_timestreams), as the reference does (:157).
    static uint32_t parallel_load_and_link(PhysicsDynamo& source, PhysicsDynamo& parallel, const uint8_t* has_future = nullptr)
    {
        uint32_t n = 0;
        check(icx_world_clone(source.world(), parallel.world(), has_future, &n), parallel.world(), "icx_world_clone");
        return n;
    }
    // ParallelCosmosContext::extract_entity_updates (:171-423): every temporal entity of `parallel` overwrites the
    // entity of `destination` whose id has the chain-link decremented; *missing = ids the destination has to
    // register first (:241-245).  Returns the number of entities written.
    static uint32_t parallel_extract(PhysicsDynamo& parallel, PhysicsDynamo& destination, uint32_t* missing = nullptr, bool register_missing = false)
    {
        uint32_t n = 0;
        check(icx_world_extract_ex(parallel.world(), destination.world(), register_missing ? ICX_EXTRACT_REGISTER : 0u, &n, missing), destination.world(), "icx_world_extract");
        return n;
    }
    IcxTimeline* timeline() { return m_tl; }

private:
    void ck(int rc) { if (rc != ICX_OK) throw std::runtime_error(std::string(icx_status_string(rc)) + " -- " + icx_timeline_last_error(m_tl)); }
    IcxTimeline* m_tl = nullptr;
};

} // namespace icx_host
