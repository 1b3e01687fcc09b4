"""CPU: the rule the vtab glue's incremental update rests on (integration/muninn_vtab_glue.c), checked against
the UNMODIFIED reference: after hnsw_insert(id) every node whose record changed is the new node or a node in its
lists (src/hnsw_algo.c:575-645: back-links, pruned or not, go to nodes the new one selected); after
hnsw_delete(id) it is the deleted node or a node in the lists the tombstone keeps (:717-805). Levels never
change, tombstones never come back, and only the newest node or a deletion moves the entry point."""
import numpy as np
import pytest

from tests.conftest import uniform


def record(g, s):
    first = int(g.first_list()[s])
    return (int(g.levels[s]), int(g.deleted[s]),
            tuple(tuple(int(x) for x in g.ids[g.targets[g.list_offsets[first + l]:g.list_offsets[first + l + 1]]])
                  for l in range(int(g.levels[s]) + 1)))


def neighbours_by_id(g, s):
    out = set()
    for lst in record(g, s)[2]:
        out.update(lst)
    return out


@pytest.mark.parametrize("metric,M,dim", [(0, 4, 8), (1, 8, 16), (2, 16, 24)])
def test_insert_and_delete_touch_only_the_node_and_its_lists(reference, metric, M, dim):
    n0, extra = 160, 36
    base = uniform(300 + dim, n0 + extra, dim)
    ref = reference.build(base[:n0], ids=np.arange(n0, dtype=np.int64) * 2 + 1, metric=metric, M=M, efc=40)
    g = ref.export()
    rng = np.random.default_rng(3)
    next_id = int(g.ids[-1]) + 1
    for step in range(extra):
        if step % 3 == 2:  # delete a live node
            live = g.ids[g.deleted == 0]
            rid = int(rng.choice(live))
            assert ref.delete(rid) == 0
            g2 = ref.export()
            s = int(np.nonzero(g2.ids == rid)[0][0])
            allowed = {rid} | neighbours_by_id(g2, s) | neighbours_by_id(g, s)
            assert g2.n == g.n
        else:  # insert above every stored id
            rid = next_id
            next_id += int(rng.integers(1, 5))
            assert ref.insert(np.array([rid], np.int64), base[n0 + step:n0 + step + 1]) == 0
            g2 = ref.export()
            assert g2.n == g.n + 1 and int(g2.ids[-1]) == rid  # appended: slots of stored nodes do not move
            allowed = {rid} | neighbours_by_id(g2, g2.n - 1)
        changed = {int(g2.ids[t]) for t in range(g.n) if record(g, t) != record(g2, t)}
        assert changed <= allowed, (step, sorted(changed - allowed))
        for t in range(g.n):
            assert g2.levels[t] == g.levels[t] and g2.deleted[t] >= g.deleted[t]
        g = g2
