#!/usr/bin/env python
"""Offline study of tile orders for frames streamed to a host canvas: list scheduling of the measured tile costs
(gpurun_out/costs_<scene>.npy, tools/scaling_probe.py RTC_DUMP_COSTS) on W warps, row completion times, and a link
that moves finished rows at a fixed rate.  usage: python tools/order_sim.py [scene]"""
import heapq
import sys

import numpy as np

W, TX, TY = 2368, 240, 270
LINK_GBS, ROW_MB = 52.0, 1920 * 4 * 24 / 1e6


def simulate(cost_us, order, speed=1.0):
    """-> kernel end (us), per-row completion times."""
    free = [0.0] * W
    heapq.heapify(free)
    done_row = np.zeros(TY)
    for t in order:
        start = heapq.heappop(free)
        end = start + cost_us[t] * speed
        heapq.heappush(free, end)
        r = t // TX
        if end > done_row[r]:
            done_row[r] = end
    return max(free), done_row


def link_end(done_row, in_order=False):
    """Finished rows go over the link one after the other (completion order, or address order if in_order)."""
    t = 0.0
    per_row = ROW_MB / LINK_GBS * 1e3  # us
    times = done_row if in_order else np.sort(done_row)
    if in_order:
        times = np.maximum.accumulate(times)
    for d in times:
        t = max(t, d) + per_row
    return t


def main():
    scene = sys.argv[1] if len(sys.argv) > 1 else "dragons"
    c = np.load("gpurun_out/costs_%s.npy" % scene).astype(np.float64) * 64 / 1965.0  # us of warp time
    n = len(c)
    total, mean = c.sum(), c.mean()
    rows = np.arange(n) // TX
    # costs were measured with the machine full; calibrate the speed so that the LPT order takes what the kernel takes
    def lpt4():
        b = np.where(c >= 8 * mean, 0, np.where(c >= 4 * mean, 1, np.where(c >= 2 * mean, 2, 3)))
        return np.argsort(b, kind="stable")
    orders = {"memory": np.arange(n), "lpt4": lpt4(),
              "heavy8_first": np.argsort(np.where(c >= 8 * mean, 0, 1), kind="stable")}
    for lead in (0.5, 1.0, 2.0):
        ahead = c * TY * W * lead / total
        orders["deadline x%.1f" % lead] = np.argsort(np.maximum(rows - ahead, 0).astype(np.int64), kind="stable")
    # deadline in WORK: row r is due when the work of the rows before it is done; start ahead by the tile's own cost
    work_before = np.concatenate([[0], np.cumsum(np.bincount(rows, weights=c, minlength=TY))])[:-1] / W
    for lead in (1.0, 2.0):
        orders["work-deadline x%.1f" % lead] = np.argsort(work_before[rows] - c * lead, kind="stable")
    # linear deadline in time over the LINK time, start ahead by own cost
    for L in (874.0, 1000.0):
        due = (rows + 1) / TY * L
        orders["link-deadline L=%d" % L] = np.argsort(due - c, kind="stable")
        orders["link-deadline L=%d x2" % L] = np.argsort(due - 2 * c, kind="stable")
    print("%s: %d tiles, %.2f s of warp time, mean %.1f us, max %.0f us, ideal %.0f us on %d warps" % (scene, n, total / 1e6, mean, c.max(), total / W, W))
    for name, o in orders.items():
        end, dr = simulate(c, o)
        q = np.percentile(dr, [10, 25, 50, 75, 90])
        print("%-24s kernel %5.0f us  rows done 10/25/50/75/90%%: %4.0f %4.0f %4.0f %4.0f %4.0f  e2e (any order) %5.0f  (address order) %5.0f" % (
            name, end, q[0], q[1], q[2], q[3], q[4], link_end(dr), link_end(dr, True)))


if __name__ == "__main__":
    main()


def rows_report(scene="dragons", lead=1.0):
    c = np.load("gpurun_out/costs_%s.npy" % scene).astype(np.float64) * 64 / 1965.0
    n = len(c)
    rows = np.arange(n) // TX
    ahead = c * TY * W * lead / c.sum()
    o = np.argsort(np.maximum(rows - ahead, 0).astype(np.int64), kind="stable")
    end, dr = simulate(c, o)
    work = np.bincount(rows, weights=c, minlength=TY)
    mx = np.array([c[rows == r].max() for r in range(TY)])
    for r in range(0, TY, 10):
        print("row %3d: done %4.0f us  work %6.0f us (cum %4.1f%%)  max tile %4.0f" % (r, dr[r], work[r], 100 * work[:r + 1].sum() / work.sum(), mx[r]))


def dma_end(done_row, min_rows=1, rate=56.4, setup=3.5):
    """Copy-engine model: one copy at a time, each takes setup + bytes / rate; when free it takes the largest... the first
    run (address order) of finished rows; runs shorter than min_rows wait unless closed / all done."""
    R = len(done_row)
    state = np.zeros(R, dtype=np.int8)  # 0 not done, 1 done, 2 copied
    t = 0.0
    order = np.argsort(done_row)
    k = 0
    copied = 0
    n_copies = 0
    while copied < R:
        while k < R and done_row[order[k]] <= t:
            state[order[k]] = 1
            k += 1
        all_in = k == R
        issued = False
        a = 0
        while a < R:
            if state[a] != 1:
                a += 1
                continue
            b = a
            while b < R and state[b] == 1:
                b += 1
            closed = (a == 0 or state[a - 1] == 2) and (b == R or state[b] == 2)
            if b - a >= min_rows or closed or all_in:
                t += setup + (b - a) * ROW_MB / rate * 1e3
                state[a:b] = 2
                copied += b - a
                n_copies += 1
                issued = True
                break
            a = b
        if not issued:
            t = done_row[order[k]] if k < R else t  # wait for the next row
    return t, n_copies


def band_orders(scene="dragons"):
    c = np.load("gpurun_out/costs_%s.npy" % scene).astype(np.float64) * 64 / 1965.0
    n = len(c)
    rows = np.arange(n) // TX
    row_work = np.bincount(rows, weights=c, minlength=TY)
    slot_us = 874.0 / TY
    print("%s: cheapest units first, tiles started ahead by lead x cost" % scene)
    for G in (1, 2, 4, 8, 16):
        nb = (TY + G - 1) // G
        band = np.arange(TY) // G
        bw = np.bincount(band, weights=row_work, minlength=nb)
        rank_of_band = np.empty(nb, dtype=np.int64)
        rank_of_band[np.argsort(bw, kind="stable")] = np.arange(nb)
        # position of every row in the sequence the rows should complete in
        seq = rank_of_band[band] * G + (np.arange(TY) % G)
        for lead in (1.0, 1.5):
            key = seq[rows] * slot_us - lead * c
            o = np.argsort(key, kind="stable")
            end, dr = simulate(c, o)
            d1, n1 = dma_end(dr, 1)
            d8, n8 = dma_end(dr, 6)
            print("  G=%2d lead %.1f: kernel %4.0f us, e2e link any order %5.0f, dma min 1 row %5.0f (%d copies), min 6 rows %5.0f (%d copies)" % (
                G, lead, end, link_end(dr), d1, n1, d8, n8))
    # reference: the plain row deadline order
    for lead in (0.5, 1.0):
        ahead = c * TY * W * lead / c.sum()
        o = np.argsort(np.maximum(rows - ahead, 0).astype(np.int64), kind="stable")
        end, dr = simulate(c, o)
        d1, n1 = dma_end(dr, 1)
        d8, n8 = dma_end(dr, 6)
        print("  top-to-bottom lead %.1f: kernel %4.0f, e2e any %5.0f, dma %5.0f (%d), min6 %5.0f (%d)" % (lead, end, link_end(dr), d1, n1, d8, n8))


def share_sim(scene="dragons", n=2, quad_x=None, buckets=(8, 4, 2)):
    """One rank's dealt share (runs of 8 order entries in turn) under the bucket order; tiles >= quad_x x mean run as four
    items of 0.47 x their cost each (measured: 790 us over four warps against 423 us for one)."""
    c = np.load("gpurun_out/costs_%s.npy" % scene).astype(np.float64) * 64 / 1965.0
    mean = c.mean()
    b = np.full(len(c), len(buckets))
    for k, x in reversed(list(enumerate(buckets))):
        b[c >= x * mean] = k
    order = np.argsort(b, kind="stable")
    mine = order[(np.arange(len(order)) // 8) % n == 0]
    items = []
    for t in mine:
        if quad_x is not None and c[t] >= quad_x * mean:
            items += [0.47 * c[t]] * 4
        else:
            items.append(c[t])
    free = [0.0] * W
    heapq.heapify(free)
    for d in items:
        s = heapq.heappop(free)
        heapq.heappush(free, s + d)
    return max(free), sum(items) / W


if __name__ == "__main__" and len(sys.argv) > 2 and sys.argv[2] == "shares":
    for n in (2, 4, 8):
        base = share_sim(sys.argv[1], n)
        print("1/%d share: one lane per pixel %4.0f us (ideal %4.0f)" % (n, base[0], base[1]), end="")
        for bk, q in (((8, 4, 2), 8), ((12, 8, 4, 2), 12), ((16, 12, 8, 4, 2), 12), ((12, 8, 4, 2), 8), ((8, 4, 2), 4)):
            r = share_sim(sys.argv[1], n, q, bk)
            print("  | buckets %s quad >= %dx: %4.0f (ideal %4.0f)" % ("/".join(map(str, bk)), q, r[0], r[1]), end="")
        print()
