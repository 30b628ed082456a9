"""The extender's HTTP bodies: ExtenderArgs -> HostPriorityList for `sort`,
ExtenderBindingArgs -> ExtenderBindingResult for `bind`.  CPU tests: the JSON codec against Python's json module.
GPU tests: the two handlers on JSON bodies against the oracle's scores and choices."""
import json
import random

import numpy as np
import pytest

RES = "aliyun.com/gpu"


def pod_body(gpus, uid="uid-1", names=None, tagged=False, extra=None):
    """An ExtenderArgs body as kube-scheduler sends it; `tagged`: the v1 json tags (lower case) instead of Go field names."""
    containers = [{"name": "c%d" % i, "image": "x", "resources": {"limits": {RES: str(g), "cpu": "500m"},
                                                                 "requests": {RES: str(g), "memory": "1Gi"}}}
                  for i, g in enumerate(gpus)]
    pod = {"metadata": {"name": "job-0", "namespace": "default", "uid": uid, "labels": {"a": "b"}},
           "spec": {"containers": containers, "nodeName": ""}, "status": {}}
    body = {"pod" if tagged else "Pod": pod, "nodes" if tagged else "Nodes": None}
    if names is not None:
        body["nodenames" if tagged else "NodeNames"] = names
    if extra:
        body.update(extra)
    return json.dumps(body)


# ------------------------------------------------------------------ CPU: codec
def test_extender_args_fields(gt):
    assert gt.wire_parse_extender_args(pod_body([2], names=["n1", "n2"])) == (2, "uid-1", ["n1", "n2"])
    assert gt.wire_parse_extender_args(pod_body([2, 1, 4], tagged=True, names=[])) == (7, "uid-1", [])
    assert gt.wire_parse_extender_args(pod_body([1])) == (1, "uid-1", None)          # no NodeNames: every node
    assert gt.wire_parse_extender_args(pod_body([], extra={"Unknown": [1, {"x": [True, None]}]})) == (0, "uid-1", None)
    # limits win over requests; a container without a limit counts its request; numbers as well as strings
    body = {"Pod": {"metadata": {"uid": "u"}, "spec": {"containers": [
        {"resources": {"limits": {RES: 3}, "requests": {RES: "1"}}},
        {"resources": {"requests": {RES: "2.0"}}},
        {"resources": {}}, {"resources": None}, {}]}}}
    assert gt.wire_parse_extender_args(json.dumps(body)) == (5, "u", None)
    # escapes and non-ASCII in names survive the round trip
    odd = ["node \"q\"", "tab\there", "snow☃man", "emoji\U0001F600", "back\\slash"]
    got = gt.wire_parse_extender_args(pod_body([1], names=odd))
    assert got[2] == odd
    assert got == gt.wire_parse_extender_args(json.dumps(json.loads(pod_body([1], names=odd)), ensure_ascii=False))


@pytest.mark.parametrize("qty", ["500m", "0.5", "-1", "1e3", "two", "", "1Gi", "1.5"])
def test_fractional_or_malformed_gpu_quantities_are_rejected(gt, qty):
    """Gaia's Fragment case (0 < m < 1) is not served (gaia.md:36); anything but a whole number is a parse error."""
    body = {"Pod": {"metadata": {"uid": "u"}, "spec": {"containers": [{"resources": {"limits": {RES: qty}}}]}}}
    with pytest.raises(gt.GtError) as e:
        gt.wire_parse_extender_args(json.dumps(body))
    assert e.value.code == -6


@pytest.mark.parametrize("bad", ["", "[]", "{", '{"Pod":', '{"Pod":{}} x', '{"NodeNames":[1]}', '{"NodeNames":{"a":1}}',
                                 '{"Pod":{"metadata":{"uid":7}}}', '{"Pod":{"spec":{"containers":[3]}}}',
                                 '{"a":"\\x"}', '{"a":"\\ud800"}', '{"a":"unterminated}', "{'a':1}",
                                 '{"a":1,}', "[" * 100 + "]" * 100])
def test_malformed_bodies_are_parse_errors(gt, bad):
    with pytest.raises(gt.GtError) as e:
        gt.wire_parse_extender_args(bad)
    assert e.value.code == -6


def test_host_priorities_and_binding_bodies(gt):
    names = ["n1", "node \"q\"", "snow☃", "ctl\x01"]
    text = gt.wire_host_priorities(names, [0, 10, 7, 3])
    assert json.loads(text) == [{"Host": n, "Score": s} for n, s in zip(names, [0, 10, 7, 3])]
    assert gt.wire_host_priorities([], []) == "[]"
    b = {"PodName": "job-0", "PodNamespace": "default", "PodUID": "uid-9", "Node": "n\"1"}
    assert gt.wire_parse_binding_args(json.dumps(b)) == b
    assert gt.wire_parse_binding_args(json.dumps({k.lower(): v for k, v in b.items()})) == b
    assert gt.wire_parse_binding_args("{}") == dict.fromkeys(b, "")
    assert json.loads(gt.wire_binding_result("")) == {"Error": ""}
    assert json.loads(gt.wire_binding_result('no "room"\n')) == {"Error": 'no "room"\n'}
    with pytest.raises(gt.GtError):
        gt.wire_parse_binding_args('{"PodName":3}')


def test_codec_fuzz_against_python_json(gt):
    """Random documents around a valid ExtenderArgs: whatever Python's json accepts with the right types parses to the
    same fields; truncations never crash and never parse."""
    rng = random.Random(11)

    def junk(d=0):
        t = rng.randrange(7 if d < 3 else 5)
        if t == 0: return rng.randrange(-5, 1000)
        if t == 1: return rng.random() * 100
        if t == 2: return rng.choice([True, False, None])
        if t in (3, 4): return "".join(rng.choice("ab \"\\/\n\té☃{}[],:") for _ in range(rng.randrange(8)))
        if t == 5: return [junk(d + 1) for _ in range(rng.randrange(4))]
        return {"k%d" % i: junk(d + 1) for i in range(rng.randrange(4))}

    for it in range(300):
        gpus = [rng.randrange(0, 5) for _ in range(rng.randrange(1, 4))]
        names = ["n%d" % rng.randrange(100) for _ in range(rng.randrange(0, 6))] if rng.random() < 0.7 else None
        extra = {"x%d" % i: junk() for i in range(rng.randrange(3))}
        body = pod_body(gpus, uid="u%d" % it, names=names, tagged=rng.random() < 0.5, extra=extra)
        if rng.random() < 0.5:
            body = json.dumps(json.loads(body), indent=rng.choice([None, 1, 3]), ensure_ascii=rng.random() < 0.5)
        assert gt.wire_parse_extender_args(body) == (sum(gpus), "u%d" % it, names)
        cut = body[:rng.randrange(len(body))]
        with pytest.raises(gt.GtError):
            gt.wire_parse_extender_args(cut)


# ------------------------------------------------------------------ GPU: the handlers on JSON bodies
@pytest.mark.gpu
def test_sort_and_bind_handlers_on_json_bodies(gt, oracle):
    lut = oracle.lut(0)
    cmax = oracle.cmax(lut)
    N = 300
    lv, ng, fr, _ = oracle.synth_cluster(4, 0xE47, 0, N)
    names = ["node-%03d" % i for i in range(N)]
    free = fr.copy()
    with gt.Allocator(N) as a:
        a.load_nodes(lv, ng, fr)
        http = gt.ExtenderHTTP(a, names)
        rng = np.random.RandomState(3)
        bound = {}
        for it in range(25):
            m = int(rng.randint(1, 9))
            uid = "pod-%d" % it
            subset = None if it % 3 == 0 else [names[i] for i in rng.permutation(N)[: rng.randint(1, 40)]] + ["not-a-node"]
            reply = json.loads(http.handle("sort", pod_body([m], uid=uid, names=subset, tagged=it % 2 == 1)))
            best, _ = oracle.exhaustive_cluster(lv, ng, free, lut, 1)
            want_names = names if subset is None else subset
            assert [r["Host"] for r in reply] == want_names
            ln = m * (m - 1) // 2
            for r in reply:
                if r["Host"] == "not-a-node":
                    assert r["Score"] == 0
                    continue
                w = int(best[names.index(r["Host"]), m - 1])
                want = 0 if w == 0xFFFF0000 else oracle.score_int(w >> 16, ln, cmax)
                assert r["Score"] == want
            # kube-scheduler picks the best host (here: highest score, first in the reply) and calls bind
            top = max((r for r in reply if r["Host"] != "not-a-node"), key=lambda r: r["Score"])
            node = names.index(top["Host"])
            res = json.loads(http.handle("bind", json.dumps({"PodName": uid, "PodNamespace": "default", "PodUID": uid,
                                                             "Node": top["Host"]}), now=1700000000 + it))
            w = int(best[node, m - 1])
            if w == 0xFFFF0000:
                assert res["Error"] != ""
                continue
            assert res == {"Error": ""}
            mask = w & 0xFFFF
            ann = http.annotations[uid]
            assert ann["ALIYUN_COM_GPU_GROUP"] == ",".join(str(g) for g in range(16) if (mask >> g) & 1)
            assert ann["ALIYUN_COM_GPU_ASSIGNED"] == "false" and ann["ALIYUN_COM_GPU_ASSUME_TIME"] == str(1700000000 + it)
            free[node] &= ~np.uint16(mask)
            bound[uid] = (node, mask)
        # errors come back in the reply body, as the reference's bind would report them
        assert "not seen by sort" in json.loads(http.handle("bind", json.dumps({"PodUID": "ghost", "Node": names[0]})))["Error"]
        http.handle("sort", pod_body([1], uid="lost"))
        assert "unknown node" in json.loads(http.handle("bind", json.dumps({"PodUID": "lost", "Node": "nowhere"})))["Error"]
        with pytest.raises(gt.GtError):
            http.handle("sort", pod_body([0], uid="none"))          # asks for no GPU: not this extender's pod
        with pytest.raises(gt.GtError):
            http.handle("sort", "{not json")
        # deleting the pods gives the GPUs back
        for uid, (node, mask) in bound.items():
            http.pod_deleted(uid)
            free[node] |= np.uint16(mask)
        with pytest.raises(gt.GtError):
            http.pod_deleted("pod-0" if "pod-0" not in bound else "ghost")
        assert (free == fr).all()
        got = a.score_all()
        ref, _ = oracle.exhaustive_cluster(lv, ng, fr, lut, 1)
        assert (got == ref).all()
