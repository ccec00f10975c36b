// TEST INFRASTRUCTURE -- not product code (see ref_harness.cpp header).
//
// ParallelCosmosContext::load_and_link / extract_entity_updates (source/spacetime/parallel_cosmos_context.cpp:60-169,
// :171-423) cannot be compiled here: the class drags in staging / assimp / asio.  Its two ENTITY LOOPS,
// however, only talk to Cosmos and EntityTimestreamMap, which do compile.  They are lifted here around the REAL objects:
//   :107-151  load_and_link   -- real Cosmos::get_signatures_ref / serialize_entity_components / register_entity /
//                                deserialize_entity_components / get_timestream_state / set_timestream_state,
//                                real EntityTimestreamMap::entity_has_timestream, real derive_causal_chain_link
//   :206-272  extract         -- the same plus the real decrement_causal_chain_link / derive_timeslice_id
//             (the worldline-shift branch :229-316 needs m_readingSteinerEntities / create_test_projectile and is not lifted:
//              the set is empty here, as it is whenever no paradox was detected)
// so that oracle/clone_oracle.py (numpy restatement) and icx_world_clone / icx_world_extract are pinned to what the reference's
// own ECS does, including its serialisation order, its "operator[] inserts {merged, 0}" state table and the coherency stamps.
#include "../../include/icx.h"

#include "core/cosmos.h"
#include "events/event_broker.h"
#include "networking/entity_timestream_map.h"
#include "physics/transform_component.h"
#include "physics/physics_component.h"
#include "physics/collider_component.h"
#include "spacetime/timestream_state.h"

#include <algorithm>
#include <map>
#include <stdexcept>

using namespace pleep;

namespace {

struct Arrays {
    uint32_t n; const uint16_t* ids; const IcxTransform* T; const IcxPhysics* P; const uint8_t* hasP; const IcxColliderComponent* C;
    const uint8_t* tstate; const uint16_t* tcoh;
};

std::shared_ptr<Cosmos> make_cosmos(std::shared_ptr<EventBroker> broker, uint32_t host)
{
    auto c = std::make_shared<Cosmos>(broker, (TimesliceId)host);
    c->register_component<TransformComponent>();
    c->register_component<PhysicsComponent>();
    c->register_component<ColliderComponent>();
    return c;
}

bool fill(Cosmos& c, const Arrays& a, uint16_t coherency)
{
    for (uint32_t i = 0; i < a.n; i++) {
        const Entity e = a.ids[i];
        if (!c.register_entity(e)) return false;
        c.add_component(e, *reinterpret_cast<const TransformComponent*>(&a.T[i]));
        if (!a.hasP || a.hasP[i]) c.add_component(e, *reinterpret_cast<const PhysicsComponent*>(&a.P[i]));
        c.add_component(e, *reinterpret_cast<const ColliderComponent*>(&a.C[i]));
        if (a.tstate && (a.tstate[i] != 0 || (a.tcoh && a.tcoh[i] != 0))) {
            c.set_coherency(a.tcoh ? a.tcoh[i] : 0);                // set_timestream_state stamps the cosmos' coherency (cosmos.cpp:161)
            c.set_timestream_state(e, (TimestreamState)a.tstate[i]);
        }
    }
    c.set_coherency(coherency);
    return true;
}

// ascending id order (= the order every other checker and the device use)
uint32_t dump(Cosmos& c, uint32_t cap, uint16_t* ids, IcxTransform* T, IcxPhysics* P, uint8_t* hasP, IcxColliderComponent* C,
              uint8_t* tstate, uint16_t* tcoh)
{
    std::vector<Entity> order;
    for (auto& kv : c.get_signatures_ref()) order.push_back(kv.first);
    std::sort(order.begin(), order.end());
    uint32_t n = 0;
    for (Entity e : order) {
        if (n >= cap) break;
        ids[n] = e;
        T[n] = *reinterpret_cast<IcxTransform*>(&c.get_component<TransformComponent>(e));
        hasP[n] = c.has_component<PhysicsComponent>(e) ? 1 : 0;
        if (hasP[n]) P[n] = *reinterpret_cast<IcxPhysics*>(&c.get_component<PhysicsComponent>(e));
        C[n] = *reinterpret_cast<IcxColliderComponent*>(&c.get_component<ColliderComponent>(e));
        auto st = c.get_timestream_state(e);
        tstate[n] = (uint8_t)st.first; tcoh[n] = st.second;
        n++;
    }
    return n;
}

} // namespace

extern "C" {

// load_and_link, :60-169.  src arrays describe sourceCosmos (mutated: forked entities are re-merged), has_future[i] != 0 <=>
// sourceFutureTimestreams->entity_has_timestream(id).  Outputs: the parallel cosmos in ascending id order.
int icxref_load_and_link(uint32_t n, const uint16_t* ids, const IcxTransform* T, const IcxPhysics* P, const uint8_t* hasP,
                         const IcxColliderComponent* C, uint8_t* tstate, uint16_t* tcoh, uint16_t src_coherency, uint32_t src_host,
                         const uint8_t* has_future,
                         uint32_t cap, uint32_t* n_out, uint16_t* o_ids, IcxTransform* o_T, IcxPhysics* o_P, uint8_t* o_hasP,
                         IcxColliderComponent* o_C, uint8_t* o_tstate, uint16_t* o_tcoh, uint16_t* o_coherency)
{
    auto broker = std::make_shared<EventBroker>();
    auto sourceCosmos = make_cosmos(broker, src_host);
    Arrays a{n, ids, T, P, hasP, C, tstate, tcoh};
    if (!fill(*sourceCosmos, a, src_coherency)) return 1;
    auto sourceFutureTimestreams = std::make_shared<EntityTimestreamMap>();
    for (uint32_t i = 0; i < n; i++)
        if (!has_future || has_future[i]) sourceFutureTimestreams->push_to_timestream(ids[i], EventMessage(events::cosmos::ENTITY_UPDATE, src_coherency));

    // :86-91 -- "keep parallel's ID as NULL_TIMESLICEID"
    auto m_currentCosmos = make_cosmos(broker, NULL_TIMESLICEID);
    m_currentCosmos->set_coherency(sourceCosmos->get_coherency());

    // :107-151, verbatim apart from logging
    for (auto signMapIt : sourceCosmos->get_signatures_ref())
    {
        if (!sourceFutureTimestreams->entity_has_timestream(signMapIt.first) && derive_causal_chain_link(signMapIt.first) == 0)
        {
            continue;
        }

        EventMessage entityUpdate(events::cosmos::ENTITY_UPDATE);
        sourceCosmos->serialize_entity_components(signMapIt.first, signMapIt.second, entityUpdate);

        if (!m_currentCosmos->register_entity(signMapIt.first, signMapIt.second))
        {
            continue;
        }
        m_currentCosmos->deserialize_entity_components(signMapIt.first, signMapIt.second, entityUpdate, ComponentCategory::all);

        if (sourceCosmos->get_timestream_state(signMapIt.first).first == TimestreamState::forked)
        {
            m_currentCosmos->set_timestream_state(signMapIt.first, TimestreamState::forked);
            sourceCosmos->set_timestream_state(signMapIt.first, TimestreamState::merged);
        }
    }

    // the source's state table after the loop
    for (uint32_t i = 0; i < n; i++) {
        auto st = sourceCosmos->get_timestream_state(ids[i]);
        tstate[i] = (uint8_t)st.first; tcoh[i] = st.second;
    }
    *n_out = dump(*m_currentCosmos, cap, o_ids, o_T, o_P, o_hasP, o_C, o_tstate, o_tcoh);
    *o_coherency = m_currentCosmos->get_coherency();
    return 0;
}

// extract_entity_updates, :171-423 (entity loop :206-272 without the worldline-shift branch).  par_* = m_currentCosmos,
// dst_* = dstCosmos (in / out, same entity set: ids the loop would have to register are reported in *n_registered and dropped).
// Returns 2 if decrement_causal_chain_link throws (chain-link 0 in the parallel cosmos).
int icxref_extract(uint32_t n_par, const uint16_t* par_ids, const IcxTransform* par_T, const IcxPhysics* par_P, const uint8_t* par_hasP,
                   const IcxColliderComponent* par_C, const uint8_t* par_tstate, const uint16_t* par_tcoh, uint16_t coherency,
                   uint32_t n_dst, const uint16_t* dst_ids, IcxTransform* dst_T, IcxPhysics* dst_P, const uint8_t* dst_hasP,
                   IcxColliderComponent* dst_C, uint8_t* dst_tstate, uint16_t* dst_tcoh, uint32_t dst_host,
                   uint32_t* n_extracted, uint32_t* n_registered)
{
    auto broker = std::make_shared<EventBroker>();
    auto m_currentCosmos = make_cosmos(broker, NULL_TIMESLICEID);
    auto dstCosmos = make_cosmos(broker, dst_host);
    Arrays ap{n_par, par_ids, par_T, par_P, par_hasP, par_C, par_tstate, par_tcoh};
    Arrays ad{n_dst, dst_ids, dst_T, dst_P, dst_hasP, dst_C, dst_tstate, dst_tcoh};
    if (!fill(*m_currentCosmos, ap, coherency) || !fill(*dstCosmos, ad, coherency)) return 1;
    uint32_t extracted = 0, registered = 0;
    try {
        // :206-272 (+ :340-349), verbatim apart from logging and the branch named above
        for (auto signMapIt : m_currentCosmos->get_signatures_ref())
        {
            EventMessage extraction(events::cosmos::ENTITY_UPDATE);

            Entity localEntity = signMapIt.first;
            if (!decrement_causal_chain_link(localEntity))
            {
                continue;
            }

            if (derive_timeslice_id(localEntity) == NULL_TIMESLICEID)
            {
                continue;
            }
            else if (true)
            {
                m_currentCosmos->serialize_entity_components(signMapIt.first, signMapIt.second, extraction);

                if (!dstCosmos->entity_exists(localEntity))
                {
                    registered++;
                    dstCosmos->register_entity(localEntity, signMapIt.second);
                }

                dstCosmos->deserialize_entity_components(localEntity, signMapIt.second, extraction, ComponentCategory::all);

                if (derive_causal_chain_link(localEntity) > 0) dstCosmos->set_timestream_state(localEntity, m_currentCosmos->get_timestream_state(signMapIt.first).first);
                extracted++;
            }

            if (dstCosmos->get_host_id() == 0U)
            {
                dstCosmos->set_timestream_state(localEntity, TimestreamState::merged);
            }
        }
    } catch (const std::range_error&) { return 2; }
    for (uint32_t i = 0; i < n_dst; i++) {
        const Entity e = dst_ids[i];
        dst_T[i] = *reinterpret_cast<IcxTransform*>(&dstCosmos->get_component<TransformComponent>(e));
        if (!dst_hasP || dst_hasP[i]) dst_P[i] = *reinterpret_cast<IcxPhysics*>(&dstCosmos->get_component<PhysicsComponent>(e));
        dst_C[i] = *reinterpret_cast<IcxColliderComponent*>(&dstCosmos->get_component<ColliderComponent>(e));
        auto st = dstCosmos->get_timestream_state(e);
        dst_tstate[i] = (uint8_t)st.first; dst_tcoh[i] = st.second;
    }
    *n_extracted = extracted; *n_registered = registered;
    return 0;
}

} // extern "C"

// ---- wire / record format ----
// The child-timestream ENTITY_UPDATE of every entity exactly as ServerNetworkDynamo::run_relays phase 4 builds it
// (server_network_dynamo.cpp:666-677) on the REAL Cosmos / EventMessage: header + body bytes at a fixed stride.
// n_dummy_before_physics: unrelated component types registered between Transform and Physics, so that the ComponentType
// indices can be those of hard_config_cosmos.h:47-58 (Transform 0, Physics 6, Collider 7).
namespace { template <int N> struct Filler { int v[2]; }; }
extern "C" int icxref_entity_update_messages(uint32_t n, const uint16_t* ids, const IcxTransform* T, const IcxPhysics* P, const uint8_t* hasP,
                                             const IcxColliderComponent* C, uint16_t coherency, int hard_config_types, int increment_link,
                                             uint8_t* out, size_t stride)
{
    auto broker = std::make_shared<EventBroker>();
    auto cosmos = std::make_shared<Cosmos>(broker, (TimesliceId)0);
    cosmos->register_component<TransformComponent>();
    if (hard_config_types) {
        cosmos->register_component<Filler<1>>(); cosmos->register_component<Filler<2>>(); cosmos->register_component<Filler<3>>();
        cosmos->register_component<Filler<4>>(); cosmos->register_component<Filler<5>>();
    }
    cosmos->register_component<PhysicsComponent>();
    cosmos->register_component<ColliderComponent>();
    for (uint32_t i = 0; i < n; i++) {
        const Entity e = ids[i];
        if (!cosmos->register_entity(e)) return 1;
        cosmos->add_component(e, *reinterpret_cast<const TransformComponent*>(&T[i]));
        if (!hasP || hasP[i]) cosmos->add_component(e, *reinterpret_cast<const PhysicsComponent*>(&P[i]));
        if (C) cosmos->add_component(e, *reinterpret_cast<const ColliderComponent*>(&C[i]));
    }
    cosmos->set_coherency(coherency);
    const uint16_t currentCoherency = cosmos->get_coherency();
    for (uint32_t i = 0; i < n; i++) {
        const Entity entity = ids[i];
        const Signature sign = cosmos->get_entity_signature(entity);
        // :666-676
        EventMessage childUpdateMsg(events::cosmos::ENTITY_UPDATE, currentCoherency);
        events::cosmos::ENTITY_UPDATE_params childUpdateInfo;
        std::memset(&childUpdateInfo, 0, sizeof(childUpdateInfo));                  // (padding bytes: the reference leaves them unspecified)
        childUpdateInfo.entity = entity; childUpdateInfo.sign = sign; childUpdateInfo.category = ComponentCategory::all;
        cosmos->serialize_entity_components(childUpdateInfo.entity, childUpdateInfo.sign, childUpdateMsg);
        if (increment_link) increment_causal_chain_link(childUpdateInfo.entity);
        childUpdateMsg << childUpdateInfo;
        uint8_t* o = out + (size_t)i * stride;
        std::memset(o, 0, stride);
        MessageHeader<EventId> h; std::memset(&h, 0, sizeof(h));
        h.id = childUpdateMsg.header.id; h.coherency = childUpdateMsg.header.coherency; h.size = childUpdateMsg.header.size;
        if (sizeof(h) + childUpdateMsg.body.size() > stride) return 2;
        std::memcpy(o, &h, sizeof(h));
        std::memcpy(o + sizeof(h), childUpdateMsg.body.data(), childUpdateMsg.body.size());
    }
    return (int)sizeof(MessageHeader<EventId>) == 12 && sizeof(events::cosmos::ENTITY_UPDATE_params) == 24 ? 0 : 3;
}
