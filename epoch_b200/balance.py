"""HARNESS, not product: the decision logic of housekeeping/balance.F90 restated so that the Python
driver mirror, the tests and tools/ can drive eb200_get_load / eb200_redistribute the way
balance_workload does.  In the integration this logic stays in the Fortran host (it works on a few
hundred integers) and nothing in the CUDA library depends on this file.

calculate_breaks follows balance.F90:2526-2669 statement by statement, including the way
load_max / load_min carry over between the perturbation sweeps."""
import numpy as np

PUSH_PER_FIELD = 5            # shared_data.F90:758
HUGE_I8 = 2 ** 63 - 1
HUGE_I4 = 2 ** 31 - 1


def ncell_min(png):
    """constants.F90:495"""
    return (png + 1) // 2 + 1


def global_load(local_loads, n_global, ng, allreduce=None):
    """get_load (balance.F90:2411-2464): sum the ranks' histograms, weight a particle push as
    push_per_field cell updates and add the field load of the interior cells.
    local_loads: per axis int64 arrays (1-ng : n+ng); allreduce(array) -> summed array."""
    nd = len(local_loads)
    out = []
    for a in range(nd):
        tot = np.asarray(local_loads[a], dtype=np.int64)
        if allreduce is not None:
            tot = np.asarray(allreduce(tot), dtype=np.int64)
        load = PUSH_PER_FIELD * tot
        other = 1
        for b in range(nd):
            if b != a:
                other *= int(n_global[b])
        load[ng:len(load) - ng] += other
        out.append(load)
    return out


def calculate_breaks(load, nproc, ng, ncm):
    """load: the array load(1-ng : sz+ng) of one axis; returns (mins, maxs), 1-based inclusive.
    ncm = ncell_min."""
    load = np.asarray(load, dtype=np.int64)
    sz = len(load) - 2 * ng

    def L(i):                       # load(i), Fortran index
        return int(load[i + ng - 1])

    def seg(i0, i1):                # SUM(load(i0:i1))
        return int(load[i0 + ng - 1:i1 + ng].sum()) if i1 >= i0 else 0

    mins = [1] * nproc
    maxs = [sz] * nproc             # maxs[p-1] = maxs(p)
    if nproc < 2:
        return mins, maxs
    ideal = int(np.floor(float(seg(1, sz)) / nproc + 0.5))
    proc, old, total = 0, 1, 0
    for idim in range(1, sz + 1):
        total_old = total
        total = total + L(idim)
        if total >= ideal:
            proc += 1
            if ideal - total_old < total - ideal:
                maxs[proc - 1] = idim - 1
            else:
                maxs[proc - 1] = idim
            nextra = old - maxs[proc - 1] + ncm
            if nextra > 0:
                maxs[proc - 1] += nextra
            if proc == nproc - 1:
                break
            old = maxs[proc - 1]
            total = total - ideal

    def backwards():
        o = sz
        for p in range(nproc - 1, 0, -1):
            if o - maxs[p - 1] < ncm:
                maxs[p - 1] = o - ncm
            o = maxs[p - 1]
    backwards()

    def spread():
        lmax, lmin, i0 = -1, HUGE_I8, 1
        for p in range(1, nproc + 1):
            i1 = maxs[p - 1]
            ll = seg(i0, i1)
            lmax, lmin = max(lmax, ll), min(lmin, ll)
            i0 = i1 + 1
        return lmax, lmin

    best = HUGE_I4                  # load_var_best = HUGE(1) (default integer)
    load_max, load_min = 0, 0       # (uninitialised in the reference until a trial is evaluated)
    for _ in range(1000):
        for i in range(1, nproc):
            old_maxs = maxs[i - 1]
            o = 0 if i == 1 else maxs[i - 2]
            new_maxs = old_maxs - 1 if old_maxs - o - 1 >= ng else old_maxs
            if new_maxs != old_maxs:
                maxs[i - 1] = new_maxs
                load_max, load_min = spread()
                if load_max - load_min < best:
                    break
                maxs[i - 1] = old_maxs
            old_maxs = maxs[i - 1]
            o = maxs[i]
            new_maxs = old_maxs + 1 if o - old_maxs - 1 >= ng else old_maxs
            if new_maxs != old_maxs:
                maxs[i - 1] = new_maxs
                load_max, load_min = spread()
                if load_max - load_min < best:
                    break
                maxs[i - 1] = old_maxs
        if load_max - load_min < best:
            best = load_max - load_min
        else:
            break
    backwards()
    o = 0
    for p in range(1, nproc):
        if maxs[p - 1] - o < ncm:
            maxs[p - 1] = o + ncm
        o = maxs[p - 1]
    for p in range(2, nproc + 1):
        mins[p - 1] = maxs[p - 2] + 1
    return mins, maxs


def balance_fraction(local_particles, local_cells, allreduce_max, allreduce_sum, nproc):
    """balance_frac of balance_workload (balance.F90:160-172)"""
    load_local = float(PUSH_PER_FIELD * local_particles + local_cells)
    load_max, load_sum = allreduce_max(load_local), allreduce_sum(load_local)
    load_av = load_sum / nproc
    return (load_av + np.sqrt(load_av)) / (load_max + np.sqrt(load_max))
