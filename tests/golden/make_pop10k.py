"""Tile the reference's 1k population x R so the UNMODIFIED reference can run BASELINE.json configs 2 and 5
(10k agents / 40k tourists, whose agents.json / tourists.json / touristsgroups.json are missing upstream:
/root/reference/.MISSING_LARGE_BLOBS).  SURVEY.md section 8d "Config 2" recipe: replica r gets agent ids + r*1000,
hhid / wpid / scid / instid / accomid offset per replica, tourists x R with tourid / groupid offsets, and the
per-day arriving group lists merged.  Shared venues (transport, airport, hospital, hubs, churches) are created by
the reference itself from data/cells.json and the population ratio (cells.py:741,819), so they are NOT replicated
here: they grow with the population as in a real 10k run.

TEST INFRASTRUCTURE ONLY (build container; /root/reference must exist).

    python tests/golden/make_pop10k.py /tmp/pop10k [R=10] [tourist_R=R]
writes <out>/{agents,tourists,touristsgroups,touristsgroupsdays,households,workplaces,schools,institutions,
institutiontypes,accommodations,smoothedagedist}.json in the layout sim.py:502-712 loads.
"""
from __future__ import annotations

import copy
import json
import os
import sys

SRC = "/root/reference/population/1kagents2ktourists2019"


def load(name):
    with open(os.path.join(SRC, name + ".json")) as f:
        return json.load(f)


def off(v, o):
    return None if v is None else v + o


def offl(lst, o):
    return None if lst is None else [x + o for x in lst]


def tile(out_dir, R=10, TR=None):
    TR = R if TR is None else TR
    os.makedirs(out_dir, exist_ok=True)
    agents = load("agents")
    households = load("households")
    workplaces = load("workplaces")
    schools = load("schools")
    institutions = load("institutions")
    insttypes = load("institutiontypes")
    accoms = load("accommodations")
    tourists = load("tourists")
    groups = load("touristsgroups")
    gdays = load("touristsgroupsdays")

    NA = len(agents)
    NH = max(h["hhid"] for h in households) + 1
    NW = max(w["wpid"] for w in workplaces) + 1
    NS = max(s["scid"] for s in schools) + 1
    NI = max(i["instid"] for i in institutions) + 1
    NACC = max(a["accomid"] for a in accoms) + 1
    NT = max(t["tourid"] for t in tourists) + 1
    NG = max(g["groupid"] for g in groups) + 1

    o_agents = {}
    o_hh, o_wp, o_sc, o_inst, o_acc = [], [], [], [], []
    for r in range(R):
        ao = r * NA
        for k, a in agents.items():
            b = dict(a)
            b["instid"] = off(a["instid"], r * NI)
            b["hhid"] = off(a["hhid"], r * NH)
            b["scid"] = off(a["scid"], r * NS)
            b["wpid"] = off(a["wpid"], r * NW)
            b["ltcfid"] = off(a.get("ltcfid"), r * NI)
            o_agents[str(int(k) + ao)] = b
        for h in households:
            o_hh.append({"member_uids": offl(h["member_uids"], ao), "hhid": h["hhid"] + r * NH,
                         "reference_uid": off(h["reference_uid"], ao), "reference_age": h["reference_age"]})
        for w in workplaces:
            o_wp.append({"member_uids": offl(w["member_uids"], ao), "wpid": w["wpid"] + r * NW, "indid": w["indid"],
                         "accomid": off(w["accomid"], r * NACC), "accomtypeid": w["accomtypeid"]})
        for s in schools:
            t = copy.deepcopy(s)
            t["scid"] = s["scid"] + r * NS
            for key in ("member_uids", "student_uids", "teacher_uids", "non_teaching_staff_uids"):
                t[key] = offl(s.get(key), ao)
            for c in t.get("classrooms", []):
                for key in ("member_uids", "student_uids", "teacher_uids"):
                    if key in c:
                        c[key] = offl(c[key], ao)
            o_sc.append(t)
        for i in institutions:
            o_inst.append({"member_uids": offl(i["member_uids"], ao), "ltcfid": off(i["ltcfid"], r * NI),
                           "resident_uids": offl(i["resident_uids"], ao), "staff_uids": offl(i["staff_uids"], ao),
                           "instid": i["instid"] + r * NI, "insttypeid": i["insttypeid"]})
        for a in accoms:
            # member_uids are room ids local to (accomtypeid, accomid) (sim.py:795-815): kept
            o_acc.append({"member_uids": list(a["member_uids"]), "accomid": a["accomid"] + r * NACC,
                          "accomtypeid": a["accomtypeid"], "roomsize": a["roomsize"]})
    o_insttypes = [{"member_uids": [m + r * NI for r in range(R) for m in it["member_uids"]],
                    "insttypeid": it["insttypeid"]} for it in insttypes]

    # tourists: replica q of the tourist set stays in replica (q % R)'s accommodations.  With TR > R two tourist
    # replicas would claim the same rooms, so TR <= R is required.
    assert TR <= R
    o_tour, o_groups = [], []
    o_gdays = {d["dayid"]: [] for d in gdays}
    for q in range(TR):
        to, go = q * NT, q * NG
        for t in tourists:
            o_tour.append({"tourid": t["tourid"] + to, "groupid": t["groupid"] + go, "subgroupid": t["subgroupid"],
                           "age": t["age"], "gender": t["gender"]})
        for g in groups:
            o_groups.append({"groupid": g["groupid"] + go,
                             "subgroupsmemberids": [offl(sg, to) for sg in g["subgroupsmemberids"]],
                             "accominfo": [[ai[0] + q * NACC, ai[1], ai[2]] for ai in g["accominfo"]],
                             "reftourid": g["reftourid"] + to, "arr": g["arr"], "dep": g["dep"],
                             "purpose": g["purpose"], "accomtype": g["accomtype"]})
        for d in gdays:
            o_gdays[d["dayid"]].extend(offl(d["member_uids"], go))
    o_gdays = [{"dayid": k, "member_uids": v} for k, v in o_gdays.items()]

    def dump(name, obj):
        with open(os.path.join(out_dir, name + ".json"), "w") as f:
            json.dump(obj, f)

    dump("agents", o_agents); dump("households", o_hh); dump("workplaces", o_wp); dump("schools", o_sc)
    dump("institutions", o_inst); dump("institutiontypes", o_insttypes); dump("accommodations", o_acc)
    dump("tourists", o_tour); dump("touristsgroups", o_groups); dump("touristsgroupsdays", o_gdays)
    dump("smoothedagedist", load("smoothedagedist"))
    return {"n_locals": len(o_agents), "n_tourists": len(o_tour), "groups": len(o_groups)}


if __name__ == "__main__":
    out = sys.argv[1]
    R = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    TR = int(sys.argv[3]) if len(sys.argv) > 3 else None
    print(tile(out, R, TR))
