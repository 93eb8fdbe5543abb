"""Host-side mirror of the four `util` helpers the hot path's callers use (reference simulator/util.py:88-150, 239-297).

The day step itself never calls these: their arithmetic lives in the CUDA kernels (vaccination multipliers and the
log-normal parameter conversion in csrc/k3_transmission.cu, the cell-type tables in csrc/common.cuh).  They are kept
here, with the reference's names and argument meaning, for the reference-side callers that sit above the boundary
(itinerary.py:2319-2330 calls `sample_log_normal` through `Epidemiology.schedule_test`, sim.py reads
`convert_celltype_to_simcelltype`) and as known-answer anchors in tests/ (checked against the C oracle and, on the
GPU, against the kernels' results).
"""
from __future__ import annotations

import math
import time

import numpy as np

from .epidemiologyclasses import CellType, SimCellType


def lognormal_params(mean, std):
    """(mean, std) of the distribution -> (mu, sigma) of the underlying normal (util.py:98-103)"""
    mu = math.log(mean ** 2 / math.sqrt(std ** 2 + mean ** 2))
    sigma = math.sqrt(math.log(std ** 2 / mean ** 2 + 1))
    return mu, sigma


def sample_log_normal(mean, std, size, isInt=False, recomputeparams=True):
    """util.py:88-115: draws from np.random.lognormal (legacy global stream).  recomputeparams converts the
    distribution's (mean, std) first; isInt rounds every draw with Python's round; size == 1 returns a scalar."""
    mu, sigma = lognormal_params(mean, std) if recomputeparams else (mean, std)
    draws = np.random.lognormal(mean=mu, sigma=sigma, size=size)
    out = [round(x) for x in draws] if isInt else draws
    return out[0] if size == 1 else out


def calculate_vaccination_multipliers(agents_vaccination_doses, agentid, current_day, vaccine_immunity,
                                      vaccine_asymptomatic, exponential_decay_interval, accum_time=None):
    """util.py:117-150 -> (immunity_multiplier, asymptomatic_multiplier, accum_time); (1.0, 1.0) if the agent has no
    dose.  d = days since the last dose: d == 0 -> (1 - rate); otherwise 1 - rate ** max(1, d / decay_interval)."""
    t0 = time.time() if accum_time is not None else None
    imm, asym = 1.0, 1.0
    if agentid in agents_vaccination_doses:
        d = current_day - agents_vaccination_doses[agentid][-1]
        if d == 0:
            imm, asym = 1.0 - vaccine_immunity, 1.0 - vaccine_asymptomatic
        else:
            x = max(1, d / exponential_decay_interval)
            imm, asym = 1.0 - vaccine_immunity ** x, 1.0 - vaccine_asymptomatic ** x
    if accum_time is not None:
        accum_time += time.time() - t0
    return imm, asym, accum_time


_SIM = {CellType.Household: SimCellType.Residence, CellType.Institution: SimCellType.Residence,
        CellType.Accommodation: SimCellType.Residence, CellType.Workplace: SimCellType.Workplace,
        CellType.School: SimCellType.School, CellType.Classroom: SimCellType.School}


def convert_celltype_to_simcelltype(cellid, cells=None, celltype=None):
    """util.py:239-268: Household / Institution / Accommodation -> Residence, Workplace, School / Classroom -> School,
    everything else -> Community"""
    if celltype is None:
        celltype = cells[cellid]["type"]
    return _SIM.get(CellType(int(celltype)), SimCellType.Community)


def get_all_contacts_ids_by_id(id, agents_contacts_keys, agents_ids=None, agents_degrees=None, shuffle=False):
    """util.py:274-297: partner ids of `id` over the pair keys, in key order; with agents_ids + agents_degrees only the
    partners whose remaining degree is > 0; optionally shuffled by the legacy global stream; None if there is none.
    (K1 holds this set as one bit row, adjacency & alive -- csrc/k1_contacts.cu.)"""
    restrict = agents_ids is not None and agents_degrees is not None
    out = []
    for pair in agents_contacts_keys:
        if id not in pair:
            continue
        other = pair[1] if pair[0] == id else pair[0]
        if not restrict or agents_degrees[agents_ids.index(other)] > 0:
            out.append(other)
    if not out:
        return None
    arr = np.array(out)
    if shuffle:
        np.random.shuffle(arr)
    return arr
