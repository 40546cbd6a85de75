"""Exhaustive interleaving check of the TMEM accumulator protocol of the overlapped-accumulator MXFP8 GEMM (csrc/gemm.cu, OVL).

Two 256-column accumulators lie 192 columns apart in TMEM — acc 0 = [0, 256), acc 1 = [192, 448) — and share columns [192, 256).
Tile t uses accumulator t % 2.  The MMA thread may start tile t once (a) the epilogue has drained accumulator t % 2 of tile t - 2
(`tmem_empty[acc]`, one arrival per epilogue warp) and (b) the warps that own the shared chunk of tile t - 1 have read it
(`ovl_free`).  Every epilogue warp walks its chunks so that the shared one comes first.  The kernel ran bit-identically to the
single-accumulator kernel on a B200 (profiles/r01c_gemm_ovl_check.txt); this model adds what a handful of runs cannot: EVERY
interleaving of the MMA thread and the epilogue warps, with the real mbarrier parity semantics (a parity wait also passes when the
barrier is an odd number of phases away in either direction), checked for

  * no MMA write into TMEM columns that some warp has not read yet (tile t - 1's shared chunk, all of tile t - 2),
  * no read of an accumulator before its tile's MMAs have completed,
  * no deadlock, every tile drained by every warp.

One TMEM lane quadrant is independent of the others except through the barrier arrival counts, so the model runs one or two
quadrants (two warps each: `half` 0 / 1) with the counts scaled.  Transcribed from gemm.cu: MMA issuer loop ("tiles_done",
`ovl_free` wait) and the staged epilogue ("walk_down", "owns_shared").  Test infrastructure only."""
from collections import deque

import pytest


def chunks_of(half, acc, out_fp32):
    """The chunk indices warp `half` reads for a tile in accumulator `acc`, in the order the kernel walks them, and which of them
    lie in the shared columns.  bf16 output: 4 chunks of 64 columns, fp32: 8 chunks of 32."""
    n = 8 if out_fp32 else 4
    mine = list(range(half, n, 2))
    if acc == 0:
        mine.reverse()                                   # walk down: the chunk in [192, 256) is the last one of accumulator 0
    shared = {n - 1} if not out_fp32 else {n - 2, n - 1}
    if acc == 1:
        shared = {0} if not out_fp32 else {0, 1}         # ... and the first one of accumulator 1
    return mine, shared


def parity_passes(completed, k):
    """mbarrier.try_wait.parity for "phase k has completed": true iff the barrier's current phase parity differs from k's."""
    return (completed & 1) != (k & 1)


def explore(ntiles, nquads, out_fp32, ovl_wait=True):
    halves = (0, 1)
    warps = [(q, h) for q in range(nquads) for h in halves]
    nw = len(warps)
    owners_per_tile = nw if out_fp32 else nw // 2       # init count of ovl_free: the warps that own a shared chunk
    # state: mma = (tile, pc) with pc 0 = waiting, 1 = issued (MMAs in flight); warps = tuple of (tile, pc) with
    #        pc 0 = wait tmem_full, 1..len(chunks) = next chunk to read, after the last read: release + next tile
    # barriers: completed-phase counts + pending arrivals: full[2], empty[2] (cnt, arrivals), ovl (cnt, arrivals)
    # issued = number of tiles whose MMAs were issued; done = number completed (in order)
    start = (0, 0, tuple((0, 0) for _ in warps), (0, 0), (0, 0), (0, 0), (0, 0), (0, 0))
    #        issued, done, warps, full0, full1 [only cnt used], empty0 (cnt, arr), empty1, ovl
    seen = {start}
    todo = deque([start])
    final = False
    while todo:
        st = todo.popleft()
        issued, done, ws, full0, full1, empty0, empty1, ovl = st
        full = [full0[0], full1[0]]
        empty = [list(empty0), list(empty1)]
        succ = []

        def pack(issued, done, ws, full, empty, ovl):
            return (issued, done, tuple(ws), (full[0], 0), (full[1], 0), tuple(empty[0]), tuple(empty[1]), tuple(ovl))

        # --- MMA thread: issue tile `issued` (the thread runs ahead of completion: tcgen05 ops are asynchronous, executed in order)
        if issued < ntiles:
            t = issued
            acc, use = t % 2, t // 2
            ok = parity_passes(empty[acc][0], use - 1) if use > 0 else True            # acc_phase ^ 1 with acc_phase = use & 1
            if ok and t > 0 and ovl_wait:
                ok = parity_passes(ovl[0], t - 1)
            if ok:
                # hazard check at issue: from now on the MMAs may write acc's columns at any time
                for (q, h), (wt, wpc) in zip(warps, ws):
                    for k in (t - 2, t - 1):
                        if k < 0:
                            continue
                        mine, shared = chunks_of(h, k % 2, out_fp32)
                        need = mine if k == t - 2 else [c for c in mine if c in shared]   # same accumulator: all of it; the other one: the overlap
                        for c in need:
                            read = wt > k or (wt == k and wpc >= 1 and mine.index(c) < wpc - 1)
                            assert read, f"MMA of tile {t} issued while warp q{q}h{h} (at tile {wt}, pc {wpc}) has not read chunk {c} of tile {k}"
                succ.append(pack(issued + 1, done, ws, full, empty, ovl))
        # --- MMA completion (tcgen05.commit -> tmem_full[acc]), in order
        if done < issued:
            f2 = list(full)
            f2[done % 2] += 1
            succ.append(pack(issued, done + 1, ws, f2, empty, ovl))
        # --- epilogue warps
        for i, ((q, h), (wt, wpc)) in enumerate(zip(warps, ws)):
            if wt >= ntiles:
                continue
            acc, use = wt % 2, wt // 2
            mine, shared = chunks_of(h, acc, out_fp32)
            w2 = list(ws)
            e2 = [list(empty[0]), list(empty[1])]
            o2 = list(ovl)
            if wpc == 0:
                if parity_passes(full[acc], use):
                    assert done > wt, f"warp q{q}h{h} passed tmem_full for tile {wt} but only {done} tiles have completed (parity aliasing)"
                    w2[i] = (wt, 1)
                    succ.append(pack(issued, done, w2, full, e2, o2))
                continue
            c = mine[wpc - 1]                                # read chunk c (tcgen05.ld + wait::ld)
            if wpc == 1 and c in shared:                     # n == 0 && owns_shared: the shared columns are in registers
                o2[1] += 1
                if o2[1] == owners_per_tile:
                    o2 = [o2[0] + 1, 0]
            if wpc == len(mine):                             # last chunk read: release_acc
                e2[acc][1] += 1
                if e2[acc][1] == nw:
                    e2[acc] = [e2[acc][0] + 1, 0]
                w2[i] = (wt + 1, 0)
            else:
                w2[i] = (wt, wpc + 1)
            succ.append(pack(issued, done, w2, full, e2, o2))
        if not succ:
            if issued == ntiles and done == ntiles and all(wt == ntiles for wt, _ in ws):
                final = True
            else:
                raise AssertionError(f"deadlock: issued {issued} done {done} warps {ws} empty {empty} ovl {ovl} full {full}")
        for s in succ:
            if s not in seen:
                seen.add(s)
                todo.append(s)
    assert final
    return len(seen)


@pytest.mark.parametrize("out_fp32", [False, True])
@pytest.mark.parametrize("ntiles,nquads", [(1, 1), (2, 1), (5, 1), (6, 1), (4, 2)])
def test_overlapped_accumulator_protocol(ntiles, nquads, out_fp32):
    assert explore(ntiles, nquads, out_fp32) > ntiles


def test_owner_sets_match_the_kernel_comment():
    """bf16 output: the shared 64-column chunk belongs to half 1 for accumulator 0 (its last chunk) and to half 0 for accumulator 1 (its
    first); fp32 output: each half owns one of the two shared 32-column chunks.  Owners read it FIRST."""
    for acc, owner in ((0, 1), (1, 0)):
        mine, shared = chunks_of(owner, acc, False)
        assert mine[0] in shared
        other, _ = chunks_of(1 - owner, acc, False)
        assert not (set(other) & shared)
    for acc in (0, 1):
        for h in (0, 1):
            mine, shared = chunks_of(h, acc, True)
            assert mine[0] in shared and len(set(mine) & shared) == 1


def test_the_checker_catches_a_missing_ovl_wait():
    """Without the `ovl_free` wait the MMA of tile t + 1 may overwrite the shared columns of tile t: the model must notice."""
    with pytest.raises(AssertionError, match="has not read chunk"):
        explore(3, 1, False, ovl_wait=False)
