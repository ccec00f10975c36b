// Host-side mirror of the jump bookkeeping of pleep::ParallelNetworkDynamo (a21, second half): the cached
// departure conditions, the per-trip arrival history with its paradox search, and the resolved-trip set.
//
//   reference member / handler                                      here
//   m_jumpConditions        parallel_network_dynamo.h:54            JumpHistory::cache_departure / on_jump_request
//   m_jumpRevisionHistory   parallel_network_dynamo.h:62            JumpHistory::on_arrival
//   m_resolvedTrips         parallel_network_dynamo.h:65            (internal)
//   JUMP_DEPARTURE from the timestream  parallel_network_dynamo.cpp:155-161
//   JUMP_ARRIVAL paradox search         parallel_network_dynamo.cpp:303-355
//   _jump_request_handler               parallel_network_dynamo.cpp:488-593
//   clear at cycle start                parallel_network_dynamo.cpp:393
//
// This is per-jump list work (a handful of entries per trip), so it stays on the host; the predicate is
// the same congruence the device scan evaluates per body (TimejumpConditions::operator==,
// spacetime/timejump_conditions.h:21-25: length2(a - b) <= 2.0f, evaluated in FP32 without contraction).
// The caller supplies origins, e.g. from PhysicsDynamo's mirror after a download or from icx_world_extract.
#pragma once
#include <cstdint>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace icx_host {

struct JumpConditions {
    uint32_t tripId;
    float origin[3];
};

// glm::length2(lhs - rhs) <= 2.0f: dot(d, d) accumulates x*x + y*y + z*z left to right
inline bool congruent(const JumpConditions& a, const JumpConditions& b)
{
    const volatile float dx = a.origin[0] - b.origin[0], dy = a.origin[1] - b.origin[1], dz = a.origin[2] - b.origin[2];
    const volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz;      // volatile: no FMA contraction whatever the flags
    const volatile float s = xx + yy;
    return (s + zz) <= 2.0f;
}

enum class JumpRequestKind { missing = 0, equivalent = 1, resolved = 2, divergent = 3 };

class JumpHistory {
public:
    static constexpr uint32_t kMissingTripId = 91339;       // parallel_network_dynamo.cpp:520

    // JUMP_DEPARTURE popped from the linked timestream: remember what the original history departed with.
    // unordered_map::insert keeps an existing entry (the reference's assert is compiled out in release).
    void cache_departure(uint32_t entity, uint32_t tripId, const float origin[3])
    {
        m_conditions.insert({entity, JumpConditions{tripId, {origin[0], origin[1], origin[2]}}});
    }

    // JUMP_ARRIVAL written into the parallel cosmos: true = paradox (the caller sends WORLDLINE_SHIFT for the
    // stripped entity).  A paradox is "the newest arrival of this trip differs from the previous one but is
    // congruent with an earlier one"; the trip is then marked resolved and its history restarted.
    bool on_arrival(uint32_t tripId, const float origin[3])
    {
        const JumpConditions now{0, {origin[0], origin[1], origin[2]}};
        bool paradox = false;
        std::vector<JumpConditions>& trip = m_history[tripId];
        if (!trip.empty() && !congruent(now, trip.back())) {
            for (const JumpConditions& earlier : trip) {
                if (congruent(now, earlier)) { paradox = true; break; }
            }
            if (paradox) { m_resolved.insert(tripId); m_history.erase(tripId); }
        }
        m_history[tripId].push_back(now);
        return paradox;
    }

    // JUMP_REQUEST raised by the parallel cosmos itself.  trip_out receives the id the forwarded request carries.
    // divergent = the caller stores the request until the matching arrival and raises DIVERGENCE.
    JumpRequestKind on_jump_request(uint32_t entity, const float origin[3], uint32_t* trip_out)
    {
        const JumpConditions now{0, {origin[0], origin[1], origin[2]}};
        JumpRequestKind kind;
        auto it = m_conditions.find(entity);
        if (it == m_conditions.end()) {
            *trip_out = kMissingTripId;
            return JumpRequestKind::missing;                // nothing cached, nothing to erase
        }
        const uint32_t trip = it->second.tripId;
        if (congruent(now, it->second)) kind = JumpRequestKind::equivalent;
        else if (m_resolved.count(trip)) { kind = JumpRequestKind::resolved; m_resolved.erase(trip); }
        else kind = JumpRequestKind::divergent;
        *trip_out = trip;
        m_conditions.erase(it);
        return kind;
    }

    void clear_departures() { m_conditions.clear(); }
    size_t trips() const { return m_history.size(); }
    size_t arrivals(uint32_t tripId) const { auto it = m_history.find(tripId); return it == m_history.end() ? 0 : it->second.size(); }
    bool is_resolved(uint32_t tripId) const { return m_resolved.count(tripId) != 0; }

private:
    std::unordered_map<uint32_t, JumpConditions> m_conditions;
    std::unordered_map<uint32_t, std::vector<JumpConditions>> m_history;
    std::unordered_set<uint32_t> m_resolved;
};

} // namespace icx_host
