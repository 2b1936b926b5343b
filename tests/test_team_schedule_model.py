"""Model of the team kernel's round scheduling (stanford_raytracer_b200/csrc/srt_team.cuh: refill, the block-wide live
count, phase transitions) in plain Python, to pin the two properties phase alignment (always on since round 2; `align=False` models the round-1 kernel) relies on:
  1. the block loop terminates and every ray of the queue is traced exactly once, knob on or off;
  2. with the knob on, every estimate round of every team falls on a block round that is a multiple of the attempt
     length (6 rounds adaptive, 4 fixed step), whatever the ray lifetimes and rejections; with it off it does not.
The arithmetic is not modelled: a ray is a list of attempt outcomes."""
import random

import pytest

IDLE, S1DX, STAGE, EST = range(4)


def run_block(rays, nteams, cyc, align):
    """rays: per ray the number of attempts it lives (accepted or rejected alike: both are `cyc` rounds)."""
    queue_head = 0
    nstages = cyc - 1  # stages 2..6 (adaptive) or 2..4 (fixed step)
    teams = [dict(phase=IDLE, exhausted=False, ray=-1, left=0, stage=0, start=False) for _ in range(nteams)]
    traced, est_rounds, sync_iter = [], [], 0
    while True:
        refill_round = (not align) or sync_iter % cyc == 0
        for t in teams:  # refill
            if t["phase"] == IDLE and not t["exhausted"] and refill_round:
                if queue_head >= len(rays):
                    t["exhausted"] = True
                else:
                    t.update(ray=queue_head, left=rays[queue_head], phase=S1DX, start=True)
                    queue_head += 1
        live = sum(1 for t in teams if t["phase"] != IDLE or (align and not t["exhausted"]))
        if live == 0:
            break
        this_round = sync_iter
        sync_iter += 1
        assert sync_iter < 10 ** 7, "block loop does not terminate"
        for t in teams:
            if t["phase"] != IDLE and t["start"]:  # top of an attempt: stop conditions
                t["start"] = False
                if t["left"] == 0:
                    traced.append(t["ray"])
                    t["phase"] = IDLE
                elif t["phase"] == STAGE:
                    t["stage"] = 2
            if t["phase"] == IDLE:
                continue
            if t["phase"] == S1DX:
                t.update(phase=STAGE, stage=2)
            elif t["phase"] == STAGE:
                if t["stage"] == nstages + 1:
                    t["phase"] = EST
                t["stage"] += 1
            else:  # EST: accept or reject, either way the next attempt starts at stage 2 in the next round
                est_rounds.append(this_round)
                t["left"] -= 1
                t.update(phase=STAGE, start=True)
    return traced, est_rounds, sync_iter


@pytest.mark.parametrize("cyc", [6, 4])
@pytest.mark.parametrize("align", [False, True])
def test_every_ray_traced_once_and_loop_terminates(cyc, align):
    rng = random.Random(cyc * 2 + align)
    rays = [rng.choice([0, 1, 2, 3, 7, 40]) for _ in range(300)]  # 0: stopped at the first check (evanescent start)
    traced, est, rounds = run_block(rays, 16, cyc, align)
    assert sorted(traced) == list(range(len(rays)))
    assert len(est) == sum(rays)
    if align:
        assert all(e % cyc == 0 for e in est)
        assert min(est) == cyc  # first wave: stencil round 0, stages in rounds 1..cyc-1, estimate in round cyc
    else:
        assert any(e % cyc != 0 for e in est)  # refills at arbitrary rounds drift out of phase


def test_alignment_costs_at_most_one_attempt_per_ray():
    rng = random.Random(7)
    rays = [rng.randint(1, 30) for _ in range(400)]
    _, _, free = run_block(rays, 16, 6, False)
    _, _, aligned = run_block(rays, 16, 6, True)
    assert aligned <= free + 6 * (len(rays) // 16 + 2)
