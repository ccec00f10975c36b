// Host-side mirror of the timestream half of pleep::TimelineApi (source/networking/timeline_api.h:53-107)
// for the dense ENTITY_UPDATE stream, forwarding to the device snapshot ring of the C ABI.
//
//   reference (per entity, per tick, heap message + 2 mutexes)      here (one call per tick per cosmos)
//   push_past_timestream(entity, msg)       timeline_api.cpp:98-108   push_past_timestream(world, coherency)
//   pop_future_timestream(entity, coh, msg) timeline_api.cpp:110-121  pop_future_timestream(world, coherency)
//   push_timestream_at_breakpoint(...)      timeline_api.cpp:123-133  push_timestream_at_breakpoint(world, coherency)
//   pop_timestream_at_breakpoint(...)       timeline_api.cpp:135-146  pop_timestream_at_breakpoint(world, coherency)
//   link_timestreams (sets breakpoints)     timeline_api.cpp:148-172  link_timestreams() / unlink_timestreams()
// Variable-length events (ENTITY_CREATED / REMOVED, JUMP_*) are rare and stay on the reference's own
// host queues; only the per-tick full-state stream moves to the GPU.
#pragma once
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
    static uint32_t parallel_extract(PhysicsDynamo& parallel, PhysicsDynamo& destination, uint32_t* missing = nullptr)
    {
        uint32_t n = 0;
        check(icx_world_extract(parallel.world(), destination.world(), &n, missing), destination.world(), "icx_world_extract");
        return n;
    }
    IcxTimeline* timeline() { return m_tl; }

private:
    void ck(int rc) { if (rc != ICX_OK) throw std::runtime_error(std::string(icx_status_string(rc)) + " -- " + icx_timeline_last_error(m_tl)); }
    IcxTimeline* m_tl = nullptr;
};

} // namespace icx_host
