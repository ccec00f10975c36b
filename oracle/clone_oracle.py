"""TEST INFRASTRUCTURE (checker, never the product path): CPU restatement in numpy of
ParallelCosmosContext::load_and_link and ::extract_entity_updates
(source/spacetime/parallel_cosmos_context.cpp:60-169, :171-423) over the AoS scene arrays.

Parity status: PINNED.  ParallelCosmosContext itself cannot be compiled in this container (staging / assimp / asio
dependencies), but its two entity loops only talk to Cosmos and EntityTimestreamMap: oracle/ref/ref_clone.cpp
lifts them around the REAL objects (real serialize / deserialize_entity_components, state table, id helpers) and
tests/test_clone_oracle.py checks this file against it on random scenes (filters, states, stamps, missing components,
chain-link 0).  Not covered: the worldline-shift ("mandela") branch :229-316.
Only tests/ may import it.
"""
import numpy as np

TS_MERGED, TS_FORKED = 0, 1
NULL_TIMESLICE, NULL_CHAINLINK = 14, 15


def _take(scene, idx):
    out = scene.copy()
    for k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh"):
        out[k] = scene[k][idx].copy()
    return out


def load_and_link(src, src_coherency, has_future=None):
    """Returns (parallel scene, parallel coherency); mutates src.tstate / src.tcoh like the reference.

    :107-151  for every source entity, in signature-map (ascending id == upload) order:
      :110    skip it if it has no future timestream and its causal chain-link is 0
      :117-137 copy all components under the same id
      :145-150 if the source state is `forked`: the copy is set forked, the source re-merged;
               Cosmos::set_timestream_state stamps the cosmos' coherency (cosmos.cpp:162)
    A fresh cosmos answers {merged, 0} for every other entity (cosmos.cpp:175-178, operator[]).
    :91       the parallel cosmos starts at the source's coherency.
    """
    n = src.n
    link = src.ids & 15
    keep = np.ones(n, bool) if has_future is None else ~((np.asarray(has_future) == 0) & (link == 0))
    idx = np.nonzero(keep)[0]
    par = _take(src, idx)
    forked = src.tstate[idx] == TS_FORKED
    par.tstate[:] = np.where(forked, TS_FORKED, TS_MERGED)
    par.tcoh[:] = np.where(forked, src_coherency, 0)
    src.tstate[idx[forked]] = TS_MERGED
    src.tcoh[idx[forked]] = src_coherency
    return par, src_coherency


def extract_entity_updates(par, dst, dst_coherency, dst_host_id, register=False):
    """Mutates dst; returns (extracted, missing).  register=True: returns (dst with the absent entities registered, extracted,
    registered) instead -- :241-245, Cosmos::register_entity leaves them {merged, 0} and the loop then treats them like the rest.

    :206-272 for every entity of the parallel cosmos:
      :212    localEntity = id with the chain-link decremented; link 15 -> ignored (:213-219),
              link 0 -> decrement_causal_chain_link throws std::range_error (ecs_types.h:148)
      :223    ids whose timeslice is NULL_TIMESLICEID were created during the run -> ignored
      :236-252 all components overwrite dst's entity localEntity (the reference registers it
              first if absent, :241-245 -- counted as `missing` here, lifecycle is host-side)
      :255    if link(localEntity) > 0 the timestream state follows (stamped with dst's coherency)
      :268-271 timeslice 0 re-merges it
    """
    if register:
        import numpy as np
        have = set(int(e) for e in dst.ids)
        absent = [i for i in range(par.n) if (int(par.ids[i]) & 15) not in (0, NULL_CHAINLINK) and (int(par.ids[i]) >> 28) != NULL_TIMESLICE
                  and int(par.ids[i]) - 1 not in have]
        if absent:
            keys = ("ids", "T", "P", "hasP", "C", "tstate", "tcoh")
            ids = np.concatenate([dst.ids, par.ids[absent] - 1]).astype(np.uint32)
            ascending = bool(np.all(np.diff(dst.ids.astype(np.int64)) > 0))
            order = np.argsort(ids, kind="stable") if ascending else np.arange(len(ids))   # icx_entities_create: sorted in, else appended
            grown = type(dst)({k: v for k, v in dst.items() if k not in keys})
            for k in keys:
                v = np.empty(dst.n + len(absent), dst[k].dtype)
                v[:dst.n] = dst[k]
                v[dst.n:] = {"ids": par.ids[absent] - 1, "tstate": TS_MERGED, "tcoh": 0}.get(k, par[k][absent] if k not in ("tstate", "tcoh") else 0)
                grown[k] = v[order]
            dst = grown
        n_ex, n_miss = extract_entity_updates(par, dst, dst_coherency, dst_host_id)
        assert n_miss == 0
        return dst, n_ex, len(absent)
    where = {int(e): j for j, e in enumerate(dst.ids)}
    extracted = missing = 0
    for i in range(par.n):
        e = int(par.ids[i])
        link = e & 15
        if link >= NULL_CHAINLINK:
            continue
        if link == 0:
            raise ValueError("chain-link 0 in the parallel cosmos (std::range_error in the reference)")
        if (e >> 28) == NULL_TIMESLICE:
            continue
        local = e - 1
        j = where.get(local)
        if j is None:
            missing += 1
            continue
        dst.T[j] = par.T[i]
        if par.hasP[i]:
            dst.P[j] = par.P[i]
        dst.C[j] = par.C[i]
        if (local & 15) > 0:
            dst.tstate[j] = par.tstate[i]
            dst.tcoh[j] = dst_coherency
        if dst_host_id == 0:
            dst.tstate[j] = TS_MERGED
            dst.tcoh[j] = dst_coherency
        extracted += 1
    return extracted, missing
