"""Toy model of the warp-vote scheduler of rt_vote.cuh (CPU only, no GPU): 32 lanes, each working through rays made
of ~31 node steps ('I'), ~1.5 leaf steps ('L') and one shading step ('S'); every iteration the warp runs ONE section
picked with the kernel's thresholds (leaf 8, shade 12, shrinking with the live lanes; bursts of 4 node steps).
Prints, for K = 1, 2, 3 rays held per lane: warp instructions per ray, thread instructions per ray, mean lanes per
section.  Used in profiles/r01_variants.md §6 to estimate what a second ray per lane could buy.

    python tools/vote_sim.py
"""
import random, sys
random.seed(1)
# per-ray script: sequence of steps 'I' (node), 'L' (leaf), then 'S'. A sample = one bounce ray (open scene), then S.
def make_ray():
    n = max(3, int(random.gauss(31, 14)))
    leaves = max(0, int(random.gauss(1.5, 1.0)))
    pos = sorted(random.sample(range(1, n), min(leaves, n - 1))) if n > 1 else []
    seq = []
    for i in range(n):
        seq.append('I')
        if i in pos: seq.append('L')
    seq.append('S')
    return seq
COST = {'I': 85, 'L': 120, 'S': 500}
OVERHEAD = 25
def run(K, rays_per_lane=200, leafT0=8, shadeT0=12, burst=4):
    # each lane holds K rays (slots); slot state = remaining script
    lanes = [[make_ray() for _ in range(K)] for _ in range(32)]
    done_rays = [0]*32
    winst = 0; tinst = 0; dens = {'I': [0,0], 'L': [0,0], 'S': [0,0]}
    while True:
        # each lane picks per mode whether it has a slot whose next step is that mode
        def lanes_with(m): return [i for i in range(32) if any(s and s[0] == m for s in lanes[i])]
        mI, mL, mS = lanes_with('I'), lanes_with('L'), lanes_with('S')
        active = len(set(mI) | set(mL) | set(mS))
        if active == 0: break
        shadeT = min(shadeT0, max(1, active*3 >> 3)); leafT = min(leafT0, max(1, active >> 2))
        winst += OVERHEAD; tinst += OVERHEAD*active
        if len(mS) >= shadeT or (not mI and not mL):
            m, reps, who = 'S', 1, mS
        elif len(mL) >= leafT or not mI:
            m, reps, who = 'L', 1, mL
        else:
            m, reps, who = 'I', burst, mI
        for r in range(reps):
            part = 0
            for i in who:
                for s in lanes[i]:
                    if s and s[0] == m:
                        s.pop(0); part += 1
                        if not s and m == 'S':
                            done_rays[i] += 1
                            if done_rays[i] + K <= rays_per_lane: s.extend(make_ray())
                        break
            if part:
                winst += COST[m]; tinst += COST[m]*part
                dens[m][0] += part; dens[m][1] += 1
    total_rays = sum(done_rays)
    return winst/total_rays, tinst/total_rays, {m: (round(v[0]/max(v[1],1),1)) for m,v in dens.items()}
for K in (1,2,3):
    print(K, run(K))
