"""Host-side BVH build + flatten producing the reference's 12-float node records (T-9).

Follows /root/reference/lib/BVH/bvhNode.js:28-73 (median split on the longest axis, stable sort
by bbox-min, one triangle per leaf), :76-93 (links), lib/BVH/bvhBuilder.js:37-54,22-26
(pre-order flatten + id fix-up) and lib/BVH/AABB.js:12-51 (triangle boxes, pad).  The reference
builds recursively; because `mid = start + floor(span/2)` depends only on counts, the same tree
comes out of a level-synchronous build, which is what lets numpy do it for 300 K triangles.
All comparisons are on float64 values exactly as the JS does (boxes are f64 until
`new Float32Array(...)` in scene.js:304).
"""
import numpy as np


def triangle_boxes(world_A, world_B, world_C):
    """AABB.bbox_triangle + pad (AABB.js:12-24, 35-51); inputs are f32-valued float64 (N,3)."""
    mn = np.minimum(world_A, np.minimum(world_B, world_C))
    mx = np.maximum(world_A, np.maximum(world_B, world_C))
    delta = 0.0001 / 2
    thin = (mx - mn) < delta
    mx = np.where(thin, mx + delta, mx)
    mn = np.where(thin, mn - delta, mn)
    return mn, mx


def build_flat_bvh(box_min, box_max):
    """Returns (nodes float32 (2N-1,12), order int64 (N,)).

    `order[k]` = index (into the input arrays) of the triangle that ends up at position k of the
    triangle buffer (scene.js:257 re-orders the triangles into BVH leaf order).
    """
    n = box_min.shape[0]
    if n == 0:
        return np.zeros((0, 12), np.float32), np.zeros((0,), np.int64)
    order = np.arange(n, dtype=np.int64)
    total = 2 * n - 1
    nodes = np.empty((total, 12), dtype=np.float64)

    # active segments of this level: start, count, pre-order node index
    seg_start = np.array([0], dtype=np.int64)
    seg_count = np.array([n], dtype=np.int64)
    seg_node = np.array([0], dtype=np.int64)

    while seg_start.size:
        bmin = box_min[order]
        bmax = box_max[order]
        # merged box of every segment (bvhNode.js:31-33, :69)
        smin = np.minimum.reduceat(bmin, seg_start, axis=0)
        smax = np.maximum.reduceat(bmax, seg_start, axis=0)
        # reduceat runs to the next start; segments are not contiguous after leaves drop out, so
        # recompute exactly over [start, start+count) where the next start is further away.
        nxt = np.append(seg_start[1:], n)
        ragged = np.nonzero(nxt != seg_start + seg_count)[0]
        for i in ragged:
            s, c = seg_start[i], seg_count[i]
            smin[i] = bmin[s:s + c].min(axis=0)
            smax[i] = bmax[s:s + c].max(axis=0)

        ext = smax - smin
        axis = np.zeros(seg_start.size, dtype=np.int64)
        axis[ext[:, 1] > ext[:, 0]] = 1
        ext_ax = ext[np.arange(seg_start.size), axis]
        axis[ext[:, 2] > ext_ax] = 2

        skip = seg_node + 2 * seg_count - 1
        skip = np.where(skip >= total, -1, skip)
        leaf = seg_count == 1
        inner = ~leaf

        nodes[seg_node, 0:3] = smin
        nodes[seg_node, 4:7] = smax
        nodes[seg_node, 10] = skip
        # leaves: bvhBuilder.js:45 -- [min, right_offset(null -> -1), max, obj.type = 2, startID, triCount, miss, axis = 0]
        ln = seg_node[leaf]
        nodes[ln, 3] = -1
        nodes[ln, 7] = 2
        nodes[ln, 8] = seg_start[leaf]
        nodes[ln, 9] = 1
        nodes[ln, 11] = 0

        if not inner.any():
            break
        s_in, c_in, node_in, ax_in = seg_start[inner], seg_count[inner], seg_node[inner], axis[inner]
        n_left = (c_in - 1) // 2 + 1           # [start .. mid], mid = start + floor(span/2)
        right_node = node_in + 2 * n_left       # left subtree holds 2*n_left - 1 nodes
        nodes[node_in, 3] = right_node
        nodes[node_in, 7] = -1
        nodes[node_in, 8] = -1
        nodes[node_in, 9] = -1
        nodes[node_in, 11] = ax_in

        # stable sort of every inner segment by bbox-min on its axis (bvhNode.js:57-61, :95-97)
        elem_seg = np.repeat(np.arange(s_in.size), c_in)
        elem_pos = np.concatenate([np.arange(s, s + c) for s, c in zip(s_in.tolist(), c_in.tolist())]) \
            if s_in.size < 64 else _ranges(s_in, c_in)
        elem_idx = order[elem_pos]
        key = box_min[elem_idx, ax_in[elem_seg]]
        perm = np.lexsort((key, elem_seg))      # stable; primary key = segment, secondary = value
        order[elem_pos] = elem_idx[perm]

        seg_start = np.stack([s_in, s_in + n_left], axis=1).reshape(-1)
        seg_count = np.stack([n_left, c_in - n_left], axis=1).reshape(-1)
        seg_node = np.stack([node_in + 1, right_node], axis=1).reshape(-1)

    return nodes.astype(np.float32), order


def _ranges(starts, counts):
    """Concatenation of arange(s, s+c) for every (s, c), vectorised."""
    total = int(counts.sum())
    offs = np.repeat(starts - np.concatenate([[0], np.cumsum(counts)[:-1]]), counts)
    return np.arange(total, dtype=np.int64) + offs


def build_flat_bvh_recursive(box_min, box_max):
    """Literal recursive restatement (small inputs only); used by tests to check the fast builder."""
    n = box_min.shape[0]
    objs = list(range(n))
    flat = []

    def gen(start, end):
        mn = box_min[objs[start:end + 1]].min(axis=0)
        mx = box_max[objs[start:end + 1]].max(axis=0)
        ext = mx - mn
        axis = 0
        if ext[1] > ext[0]:
            axis = 1
        if ext[2] > ext[axis]:
            axis = 2
        node = {"min": mn, "max": mx, "left": None, "right": None, "axis": 0, "start": -1}
        if end - start <= 0:
            node["start"] = start
        else:
            sub = objs[start:end + 1]
            sub.sort(key=lambda i: box_min[i, axis])      # Python's sort is stable, like JS's
            objs[start:end + 1] = sub
            mid = start + (end - start) // 2
            node["left"] = gen(start, mid)
            node["right"] = gen(mid + 1, end)
            node["axis"] = axis
        return node

    def links(node, next_right):
        node["miss"] = next_right
        if node["left"] is not None:
            links(node["left"], node["right"])
            links(node["right"], next_right)

    def flatten(node):
        node["id"] = len(flat)
        flat.append(node)
        if node["left"] is not None:
            flatten(node["left"])
            flatten(node["right"])

    if n == 0:
        return np.zeros((0, 12), np.float32), np.zeros((0,), np.int64)
    root = gen(0, n - 1)
    links(root, None)
    flatten(root)
    out = np.empty((len(flat), 12), dtype=np.float64)
    for i, nd in enumerate(flat):
        miss = nd["miss"]["id"] if nd["miss"] is not None else -1
        if nd["left"] is None:
            out[i] = [*nd["min"], -1, *nd["max"], 2, nd["start"], 1, miss, 0]
        else:
            out[i] = [*nd["min"], nd["right"]["id"], *nd["max"], -1, -1, -1, miss, nd["axis"]]
    return out.astype(np.float32), np.array(objs, dtype=np.int64)
