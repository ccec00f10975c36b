// TEST INFRASTRUCTURE -- not product code (see ref_harness.cpp header).
//
// The reference's own timestream containers and predicates, driven headlessly:
//   EntityTimestreamMap / TsBreakpointQueue   source/networking/entity_timestream_map.h, ts_breakpoint_queue.h
//   Message<EventId> (LIFO byte stack)        source/events/message.h
//   coherency_greater_or_equal                source/events/message.h:179-183
//   is_divergent                              source/spacetime/timestream_state.h:41-44
//   TimejumpConditions::operator==            source/spacetime/timejump_conditions.h:21-25
// The dynamo code that calls them (ServerNetworkDynamo / ParallelNetworkDynamo::run_relays) does not
// compile here (asio is not vendored), so the record / replay loops below restate
//   record : server_network_dynamo.cpp:642-679   parallel_network_dynamo.cpp:369-385
//   replay : server_network_dynamo.cpp:377-421   parallel_network_dynamo.cpp:51-89
//   ageing : server_network_dynamo.cpp:13-15, :681-705
//   jump bookkeeping (cached departures, paradox search, request classification):
//            parallel_network_dynamo.cpp:155-161, :303-355, :488-593 -- the three containers are declared as in
//            parallel_network_dynamo.h:54-65 and compared with the REAL TimejumpConditions::operator==
// around the REAL containers, with the entity's components serialised into a real EventMessage in
// the reference's order (descending component type so that they pop ascending,
// ecs/component_registry.h:238-253: Transform first registered, then Physics).
#include "../../include/icx.h"

#include "events/event_types.h"
#include "events/message.h"
#include "networking/entity_timestream_map.h"
#include "physics/transform_component.h"
#include "physics/physics_component.h"
#include "physics/collider_component.h"
#include "spacetime/timestream_state.h"
// timejump_conditions.h pulls in Cosmos only for helpers that are not used here
#include "core/cosmos.h"
#include "spacetime/timejump_conditions.h"

#include <algorithm>
#include <cstring>
#include <memory>
#include <unordered_map>
#include <unordered_set>
#include <vector>

using namespace pleep;

namespace {

struct RefTimeline {
    EntityTimestreamMap map;
    EntityTimestreamMap events;     // ENTITY_CREATED / REMOVED / JUMP_* messages: the same real container, keyed by the same ids
};

inline IcxEntity widen(Entity e) { return e == NULL_ENTITY ? ICX_NULL_ENTITY : (((IcxEntity)(e >> 12) << 28) | ((IcxEntity)((e >> 4) & 0xFFu) << 4) | (e & 0xFu)); }

inline Entity narrow(IcxEntity e) { return (Entity)(((e >> 28) << 12) | (((e >> 4) & 0xFFu) << 4) | (e & 0xFu)); }

} // namespace

extern "C" {

void* icxref_tl_create(void) { return new RefTimeline(); }
void icxref_tl_destroy(void* h) { delete static_cast<RefTimeline*>(h); }

// link_timestreams -> set breakpoints on every stream (entity_timestream_map.h:168-199)
void icxref_tl_breakpoint(void* h, int set)
{
    RefTimeline* tl = static_cast<RefTimeline*>(h);
    if (set) { tl->map.set_breakpoints(); tl->events.set_breakpoints(); } else { tl->map.remove_breakpoints(); tl->events.remove_breakpoints(); }
}

// Event records as the handlers push them (server_network_dynamo.cpp:748-811, :878-976): REAL EventMessages with the
// reference's own params structs in the real per-entity streams.  ev->entity / ev->source are already the receiver's ids.
int icxref_tl_event_push(void* h, uint16_t coherency, const IcxTimelineEvent* ev, uint32_t count, int at_breakpoint)
{
    RefTimeline* tl = static_cast<RefTimeline*>(h);
    for (uint32_t i = 0; i < count; i++) {
        const Entity e = narrow(ev[i].entity), src = ev[i].source == ICX_NULL_ENTITY ? NULL_ENTITY : narrow(ev[i].source);
        EventMessage msg(1, coherency);
        switch (ev[i].type) {
        case ICX_EVENT_ENTITY_CREATED: { msg.header.id = events::cosmos::ENTITY_CREATED; events::cosmos::ENTITY_CREATED_params p; p.entity = e; p.sign = Signature(ev[i].payload[0]); p.source = src; msg << p; } break;
        case ICX_EVENT_ENTITY_REMOVED: { msg.header.id = events::cosmos::ENTITY_REMOVED; events::cosmos::ENTITY_REMOVED_params p; p.entity = e; p.source = src; msg << p; } break;
        default: {
            msg.header.id = ev[i].type == ICX_EVENT_JUMP_REQUEST ? events::network::JUMP_REQUEST
                          : ev[i].type == ICX_EVENT_JUMP_DEPARTURE ? events::network::JUMP_DEPARTURE : events::network::JUMP_ARRIVAL;
            events::network::JUMP_params p; p.entity = e; p.tripId = ev[i].payload[0]; p.timesliceDelta = (int)ev[i].payload[1]; p.isFocal = ev[i].payload[2] != 0;
            p.sign = Signature(ev[i].payload[3]); msg << p; } break;
        }
        if (at_breakpoint) tl->events.push_to_timestream_at_breakpoint(e, msg); else tl->events.push_to_timestream(e, msg);
    }
    return 0;
}

// The pop loop of server_network_dynamo.cpp:377-392 (parallel_network_dynamo.cpp:51-77 at the breakpoint) over the event
// streams: every stream, every message that is due, stale ones dropped.  Streams are visited in ascending id order here
// (the reference walks an unordered container); per stream the order is the container's.
int icxref_tl_events_pop(void* h, uint16_t now, int mode, IcxTimelineEvent* out, size_t cap, size_t* n_out)
{
    RefTimeline* tl = static_cast<RefTimeline*>(h);
    std::vector<Entity> es = tl->events.get_entities_with_streams();
    std::sort(es.begin(), es.end());
    size_t n = 0;
    for (Entity e : es) {
        EventMessage evnt(1);
        for (;;) {
            const bool got = mode == ICX_REPLAY_PARALLEL ? tl->events.pop_from_timestream_at_breakpoint(e, now, evnt) : tl->events.pop_from_timestream(e, now, evnt);
            if (!got) break;
            if (evnt.header.coherency < now) continue;                                          // :391
            IcxTimelineEvent r; std::memset(&r, 0, sizeof(r));
            r.coherency = evnt.header.coherency;
            if (evnt.header.id == events::cosmos::ENTITY_CREATED) { events::cosmos::ENTITY_CREATED_params p; evnt >> p; r.type = ICX_EVENT_ENTITY_CREATED; r.entity = widen(p.entity); r.source = widen(p.source); r.payload[0] = (uint32_t)p.sign.to_ulong(); }
            else if (evnt.header.id == events::cosmos::ENTITY_REMOVED) { events::cosmos::ENTITY_REMOVED_params p; evnt >> p; r.type = ICX_EVENT_ENTITY_REMOVED; r.entity = widen(p.entity); r.source = widen(p.source); }
            else {
                events::network::JUMP_params p; evnt >> p;
                r.type = evnt.header.id == events::network::JUMP_REQUEST ? ICX_EVENT_JUMP_REQUEST : evnt.header.id == events::network::JUMP_DEPARTURE ? ICX_EVENT_JUMP_DEPARTURE : ICX_EVENT_JUMP_ARRIVAL;
                r.entity = widen(p.entity); r.source = ICX_NULL_ENTITY; r.payload[0] = p.tripId; r.payload[1] = (uint32_t)p.timesliceDelta; r.payload[2] = p.isFocal ? 1u : 0u; r.payload[3] = (uint32_t)p.sign.to_ulong();
            }
            if (out && n < cap) out[n] = r;
            n++;
        }
    }
    if (n_out) *n_out = n;
    return 0;
}

// Record every entity at `coherency`.  link_inc: the server increments the causal chain-link of the
// id it pushes into the past (server_network_dynamo.cpp:673); entities are keyed by that id.
int icxref_tl_record(void* h, uint32_t n, const IcxEntity* ids, const IcxTransform* T, const IcxPhysics* P, const IcxColliderComponent* C,
                     uint16_t coherency, int at_breakpoint, int link_inc)
{
    RefTimeline* tl = static_cast<RefTimeline*>(h);
    for (uint32_t i = 0; i < n; i++) {
        EventMessage msg(events::cosmos::ENTITY_UPDATE, coherency);
        // serialize_entity_components: highest component type first (hard_config_cosmos.h:47-58 registers
        // Transform = 0, ..., Physics = 6, Collider = 7)
        if (C) msg << *reinterpret_cast<const ColliderComponent*>(&C[i]);
        msg << *reinterpret_cast<const PhysicsComponent*>(&P[i]);
        msg << *reinterpret_cast<const TransformComponent*>(&T[i]);
        Entity e = narrow(ids[i]);
        if (link_inc) increment_causal_chain_link(e);
        events::cosmos::ENTITY_UPDATE_params info = { e, Signature(C ? 0xC1 : 0x41), ComponentCategory::all };
        msg << info;
        if (at_breakpoint) tl->map.push_to_timestream_at_breakpoint(e, msg);
        else tl->map.push_to_timestream(e, msg);
    }
    return 0;
}

// Replay at `now`.  ids are the RECEIVING cosmos' ids (already carrying the incremented link).
// mode ICX_REPLAY_SERVER: pop_from_timestream, drop stale, apply to non-divergent entities.
// mode ICX_REPLAY_PARALLEL: pop_from_timestream_at_breakpoint, upstream components only (= none here).
// Returns the number of entities whose components were overwritten.
int icxref_tl_replay(void* h, uint32_t n, const IcxEntity* ids, IcxTransform* T, IcxPhysics* P, IcxColliderComponent* C, const uint8_t* tstate,
                     uint16_t now, int mode)
{
    RefTimeline* tl = static_cast<RefTimeline*>(h);
    int applied = 0;
    for (uint32_t i = 0; i < n; i++) {
        const Entity e = narrow(ids[i]);
        EventMessage evnt(1);
        for (;;) {
            const bool got = mode == ICX_REPLAY_PARALLEL ? tl->map.pop_from_timestream_at_breakpoint(e, now, evnt)
                                                         : tl->map.pop_from_timestream(e, now, evnt);
            if (!got) break;
            if (mode == ICX_REPLAY_SERVER && evnt.header.coherency < now) continue;         // :391
            events::cosmos::ENTITY_UPDATE_params info;
            evnt >> info;
            if (mode == ICX_REPLAY_PARALLEL) continue;                                      // upstream only: nothing of ours
            if (!is_divergent((TimestreamState)(tstate ? tstate[i] : 0))) {                 // :408-412
                TransformComponent t; PhysicsComponent p;
                evnt >> t; evnt >> p;                                                       // ascending component type
                T[i] = *reinterpret_cast<IcxTransform*>(&t);
                P[i] = *reinterpret_cast<IcxPhysics*>(&p);
                if (info.sign.test(7)) {                                                    // deserialize_entity_components walks the signature
                    ColliderComponent c; evnt >> c;
                    if (C) std::memcpy(&C[i], &c, sizeof(IcxColliderComponent));
                }
                applied++;
            }
        }
    }
    return applied;
}

// Congruence of live vs recorded origins with the reference's own operator==.
int icxref_congruent(uint32_t n, const IcxTransform* live, const IcxTransform* rec, uint8_t* out)
{
    for (uint32_t i = 0; i < n; i++) {
        TimejumpConditions a{0, glm::vec3(live[i].origin[0], live[i].origin[1], live[i].origin[2])};
        TimejumpConditions b{0, glm::vec3(rec[i].origin[0], rec[i].origin[1], rec[i].origin[2])};
        out[i] = (a == b) ? 1 : 0;
    }
    return 0;
}

int icxref_coherency_ge(uint16_t a, uint16_t b) { return coherency_greater_or_equal(a, b) ? 1 : 0; }

// ---- jump bookkeeping of ParallelNetworkDynamo over the real TimejumpConditions ----
struct RefJumps {
    std::unordered_map<uint32_t, TimejumpConditions> m_jumpConditions;                       // keyed by the (wide) entity id
    std::unordered_map<uint32_t, std::vector<TimejumpConditions>> m_jumpRevisionHistory;
    std::unordered_set<uint32_t> m_resolvedTrips;
};
void* icxref_jumps_create(void) { return new RefJumps(); }
void icxref_jumps_destroy(void* h) { delete static_cast<RefJumps*>(h); }
void icxref_jumps_clear_departures(void* h) { static_cast<RefJumps*>(h)->m_jumpConditions.clear(); }      // :393

// JUMP_DEPARTURE from the timestream, :155-161
void icxref_jumps_departure(void* h, uint32_t entity, uint32_t tripId, const float* o)
{
    RefJumps& J = *static_cast<RefJumps*>(h);
    J.m_jumpConditions.insert({ entity, TimejumpConditions{ tripId, glm::vec3(o[0], o[1], o[2]) } });
}

// JUMP_ARRIVAL, :303-355; returns 1 when the reference would log PARADOX DETECTED and send WORLDLINE_SHIFT
int icxref_jumps_arrival(void* h, uint32_t tripId, const float* o)
{
    RefJumps& J = *static_cast<RefJumps*>(h);
    TimejumpConditions newConditions{ 0, glm::vec3(o[0], o[1], o[2]) };
    int paradox = 0;
    if (!J.m_jumpRevisionHistory[tripId].empty() && !(newConditions == J.m_jumpRevisionHistory[tripId].back()))
    {
        for (TimejumpConditions& cond : J.m_jumpRevisionHistory[tripId])
        {
            if (newConditions == cond)
            {
                paradox = 1;
                J.m_resolvedTrips.insert(tripId);
                J.m_jumpRevisionHistory.erase(tripId);
                break;
            }
        }
    }
    J.m_jumpRevisionHistory[tripId].push_back(newConditions);
    return paradox;
}

// _jump_request_handler, :488-593; returns 0 missing / 1 equivalent / 2 resolved / 3 divergent
int icxref_jumps_request(void* h, uint32_t entity, const float* o, uint32_t* tripOut)
{
    RefJumps& J = *static_cast<RefJumps*>(h);
    TimejumpConditions newConditions{ 0, glm::vec3(o[0], o[1], o[2]) };
    int kind;
    if (J.m_jumpConditions.count(entity) < 1) { *tripOut = 91339; kind = 0; }
    else if (newConditions == J.m_jumpConditions[entity]) { *tripOut = J.m_jumpConditions[entity].tripId; kind = 1; }
    else if (J.m_resolvedTrips.count(J.m_jumpConditions[entity].tripId))
    {
        *tripOut = J.m_jumpConditions[entity].tripId;
        J.m_resolvedTrips.erase(J.m_jumpConditions[entity].tripId);
        kind = 2;
    }
    else { *tripOut = J.m_jumpConditions[entity].tripId; kind = 3; }
    J.m_jumpConditions.erase(entity);
    return kind;
}

} // extern "C"
