"""Data-plane life cycle at the launcher's rendezvous store (launcher/rendezvous.py): job-uid
scoped keys, gang generations across replica restarts, descriptor passing, clean-up on job
deletion. Behavioural reference: a retryable replica is re-created under the same name
(pkg/controller.v1/tensorflow/pod.go:140-159) and the e2e suite re-creates a whole job under the
same name (py/kubeflow/tf_operator/simple_tfjob_tests.py:26-90, --num_trials 2)."""
import os
import threading
import time

import pytest

from tf_operator_b200.launcher.rendezvous import RendezvousClient, RendezvousServer


@pytest.fixture()
def srv():
    s = RendezvousServer().start()
    yield s
    s.stop()


def _run(fns):
    out, errs = [None] * len(fns), []

    def call(i, fn):
        try:
            out[i] = fn()
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=call, args=(i, f)) for i, f in enumerate(fns)]
    [t.start() for t in ts]
    [t.join(20) for t in ts]
    assert not errs, errs
    return out


def test_first_gang_forms_generation_zero_and_exchanges_handles(srv):
    scope = "ns/job/uid-1/comm0"

    def rank(r):
        c = RendezvousClient(srv.addr)
        g = c.join(scope, r, 3)
        vals = c.all_gather(f"{scope}/g{g}/ipc", r, 3, f"handle-{r}".encode())
        c.barrier(f"{scope}/g{g}/ready", 3)
        c.close()
        return g, vals
    res = _run([lambda r=r: rank(r) for r in range(3)])
    assert all(g == 0 for g, _ in res)
    assert all(v == [b"handle-0", b"handle-1", b"handle-2"] for _, v in res)


def test_restarted_rank_opens_a_new_generation_and_stale_keys_are_gone(srv):
    scope = "ns/job/uid-1/comm0"
    _run([lambda r=r: RendezvousClient(srv.addr).join(scope, r, 2) for r in range(2)])
    c = RendezvousClient(srv.addr)
    c.set(f"{scope}/g0/ipc/0", b"old-0")
    c.set(f"{scope}/g0/ipc/1", b"old-1")
    c.add(f"{scope}/g0/ready/count", 2)
    # rank 1 is re-created by the operator: it must NOT see rank 0's old handle (freed memory)
    joined = {}

    def rejoin_1():
        cli = RendezvousClient(srv.addr)
        g = cli.join(scope, 1, 2, timeout=10)
        joined[1] = g
        return cli.all_gather(f"{scope}/g{g}/ipc", 1, 2, b"new-1")
    t = threading.Thread(target=rejoin_1)
    t.start()
    time.sleep(0.3)
    assert 1 not in joined                      # waits for the rest of the gang in generation 1
    assert srv.keys(f"{scope}/g0/") == []       # generation 0 dropped the moment it was superseded
    with pytest.raises(TimeoutError):
        c.get(f"{scope}/g0/ipc/0", timeout=0.2)
    # rank 0 times out in its collective, exits 138, is re-created and joins too
    cli0 = RendezvousClient(srv.addr)
    g0 = cli0.join(scope, 0, 2, timeout=10)
    vals0 = cli0.all_gather(f"{scope}/g{g0}/ipc", 0, 2, b"new-0")
    t.join(10)
    assert g0 == joined[1] == 1 and vals0 == [b"new-0", b"new-1"]


def test_waiter_of_a_superseded_generation_rejoins_by_itself(srv):
    scope = "ns/job/uid-2/comm0"
    _run([lambda r=r: RendezvousClient(srv.addr).join(scope, r, 3) for r in range(3)])
    from tf_operator_b200.launcher.rendezvous import Replaced
    got, zombie = {}, []

    def join(r, delay, who):
        time.sleep(delay)
        try:
            got[who] = RendezvousClient(srv.addr).join(scope, r, 3, timeout=15)
        except Replaced as e:
            zombie.append((who, e))
    # rank 0 restarts first (-> generation 1) and waits. Then rank 2 restarts twice in a row: its
    # second incarnation opens generation 2. Rank 0's pending join is superseded and re-joins by
    # itself; rank 2's first incarnation learns it has been replaced and gives up; rank 1 follows.
    ts = [threading.Thread(target=join, args=(0, 0.0, "r0")),
          threading.Thread(target=join, args=(2, 0.2, "r2-first")),
          threading.Thread(target=join, args=(2, 0.5, "r2-second")),
          threading.Thread(target=join, args=(1, 0.8, "r1"))]
    [t.start() for t in ts]
    [t.join(20) for t in ts]
    assert [w for w, _ in zombie] == ["r2-first"]
    assert got == {"r0": 2, "r2-second": 2, "r1": 2}


def test_same_named_job_with_a_new_uid_never_sees_the_old_jobs_keys(srv):
    c = RendezvousClient(srv.addr)
    c.set("ns/job/uid-old/comm0/g0/ipc/0", b"stale")
    with pytest.raises(TimeoutError):
        c.get("ns/job/uid-new/comm0/g0/ipc/0", timeout=0.2)
    g = _run([lambda r=r: RendezvousClient(srv.addr).join("ns/job/uid-new/comm0", r, 2)
              for r in range(2)])
    assert g == [0, 0]


def test_descriptor_broker_passes_an_open_file_between_processes(srv, tmp_path):
    """SCM_RIGHTS through the broker: what carries the multicast (NVLS) handle from the rank that
    created it to the ranks that import it."""
    r, w = os.pipe()
    a = RendezvousClient(srv.addr, uds_path=srv.uds_path)
    a.fd_put("ns/job/uid/comm0/g0/nvls/fd", w)
    os.close(w)                                  # the broker holds its own duplicate
    pid = os.fork()
    if pid == 0:                                 # "another replica"
        try:
            b = RendezvousClient(srv.addr, uds_path=srv.uds_path)
            fd = b.fd_get("ns/job/uid/comm0/g0/nvls/fd", timeout=5)
            os.write(fd, b"through the broker")
            os.close(fd)
            os._exit(0)
        except BaseException:  # noqa: BLE001
            os._exit(1)
    _, status = os.waitpid(pid, 0)
    assert status == 0
    assert os.read(r, 64) == b"through the broker"
    os.close(r)
    with pytest.raises(TimeoutError):
        a.fd_get("ns/job/uid/comm0/g0/nvls/missing", timeout=0.2)
    assert srv.delete_prefix("ns/job/uid/") == 1
    with pytest.raises(TimeoutError):
        a.fd_get("ns/job/uid/comm0/g0/nvls/fd", timeout=0.2)


def test_deleting_the_job_object_clears_its_rendezvous_state():
    from tf_operator_b200.cluster import LocalCluster
    job = {"apiVersion": "kubeflow.org/v1", "kind": "PyTorchJob",
           "metadata": {"name": "rdv-clean", "namespace": "default"},
           "spec": {"pytorchReplicaSpecs": {"Master": {"replicas": 1, "restartPolicy": "Never", "template": {
               "spec": {"containers": [{"name": "pytorch", "image": "x",
                                        "command": ["python", "-c", "import time; time.sleep(30)"]}]}}}}}}
    with LocalCluster() as c:
        created = c.store.create("pytorchjobs", job)
        uid = created["metadata"]["uid"]
        cli = RendezvousClient(c.rendezvous.addr)
        scope = f"default/rdv-clean/{uid}/comm0"
        assert cli.join(scope, 0, 1) == 0
        cli.set(f"{scope}/g0/ipc/0", b"h")
        cli.set("default/other-job/uid-x/comm0/g0/ipc/0", b"keep")
        # the replica's environment carries the job uid the scope is built from
        deadline = time.time() + 10
        pods = []
        while time.time() < deadline and not pods:
            pods = [p for p in c.store.list("pods") if p["metadata"]["name"] == "rdv-clean-master-0"
                    and (p.get("status") or {}).get("phase") == "Running"]
            time.sleep(0.05)
        assert pods
        env = c.kubelet._env_for(c.kubelet._pods[("default", "rdv-clean-master-0")], pods[0],
                                 pods[0]["spec"]["containers"][0])
        assert env["TFOP_JOB_UID"] == uid and env["TFOP_INCARNATION"] == "0"
        assert env["TFOP_RENDEZVOUS_UDS"] == c.rendezvous.uds_path
        c.store.delete("pytorchjobs", "default", "rdv-clean")
        deadline = time.time() + 5
        while time.time() < deadline and c.rendezvous.keys(f"default/rdv-clean/{uid}/"):
            time.sleep(0.05)
        assert c.rendezvous.keys(f"default/rdv-clean/{uid}/") == []
        assert c.rendezvous.keys("default/other-job/") == ["default/other-job/uid-x/comm0/g0/ipc/0"]
        # a second job of the same name gets a different uid -> a different scope
        again = c.store.create("pytorchjobs", job)
        assert again["metadata"]["uid"] != uid


def test_a_wait_inside_a_generation_ends_when_the_generation_does(srv):
    """Seen on 8 GPUs: rank A gets through join of generation g a moment before a late restart of
    rank B supersedes g. A then waits for B's handle under g - a key nobody will ever publish (the
    generation's keys are dropped) - while everyone else waits for A in g+1; all time out. A wait
    that names its generation raises Superseded at once, and A re-joins."""
    from tf_operator_b200.launcher.rendezvous import Superseded
    scope = "ns/job/uid-9/comm0"
    a, b = RendezvousClient(srv.addr), RendezvousClient(srv.addr)
    assert _run([lambda: a.join(scope, 0, 2), lambda: b.join(scope, 1, 2)]) == [0, 0]
    got = {}

    def waits(kind):
        t0 = time.time()
        try:
            if kind == "get":
                a.get(f"{scope}/g0/ipc/1", timeout=30, scope=scope, gen=0)
            elif kind == "barrier":
                RendezvousClient(srv.addr).barrier(f"{scope}/g0/ready", 2, timeout=30, scope=scope, gen=0)
            else:
                RendezvousClient(srv.addr, uds_path=srv.uds_path).fd_get(
                    f"{scope}/g0/nvls/fd", timeout=30, scope=scope, gen=0)
            got[kind] = "returned"
        except Superseded:
            got[kind] = time.time() - t0
    ts = [threading.Thread(target=waits, args=(k,)) for k in ("get", "barrier", "fd")]
    [t.start() for t in ts]
    time.sleep(0.3)
    assert not got                                    # all three are blocked
    b2 = RendezvousClient(srv.addr)                   # rank 1 restarted: a new identity
    j = threading.Thread(target=lambda: got.setdefault("b2", b2.join(scope, 1, 2, timeout=20)))
    j.start()
    [t.join(10) for t in ts]
    assert all(isinstance(got.get(k), float) and got[k] < 5 for k in ("get", "barrier", "fd")), got
    # A goes back to join and the two meet in generation 1; an unscoped wait keeps its old meaning
    assert a.join(scope, 0, 2, timeout=20) == 1
    j.join(10)
    assert got["b2"] == 1
    with pytest.raises(TimeoutError):
        a.get(f"{scope}/g1/never", timeout=0.3)


class _FakeComm:
    """Stands in for runtime.symm.SymmetricComm on a CPU box: records what rendezvous_comm does."""
    made = []

    def __init__(self, rank, world, handles):
        self.rank, self.world, self.handles = rank, world, handles
        self.nvls_bytes = -1
        self.nvls_calls = []
        _FakeComm.made.append(self)

    @classmethod
    def create(cls, rank, world, nbytes, exchange, device=None):
        return cls(rank, world, exchange(f"h{rank}-{nbytes}".encode()))

    def enable_nvls(self, nbytes):
        cli, prefix = self._rdv
        scope, gen = self._rdv_gen
        self.nvls_calls.append((prefix, scope, gen))
        cli.barrier(prefix + "/nvls/bound", self.world, scope=scope, gen=gen)
        self.nvls_bytes = nbytes
        return True


def test_rendezvous_comm_forms_ipc_and_nvls_in_one_generation_and_rejoins_when_superseded(srv, monkeypatch):
    """runtime.env.rendezvous_comm with the communicator faked: the handles and the multicast
    set-up happen under <scope>/g<gen>/ with every wait naming (scope, gen); a restart of a peer
    in the middle sends the survivor back to join, and both end up in the next generation."""
    import tf_operator_b200.runtime.symm as symm
    from tf_operator_b200.runtime.env import _ABANDONED, rendezvous_comm
    monkeypatch.setattr(symm, "SymmetricComm", _FakeComm)
    _FakeComm.made.clear()
    scope = "ns/job/uid-7/comm0"
    comms = _run([lambda r=r: rendezvous_comm(srv.addr, srv.uds_path, scope, r, 2, 1 << 20, nvls_bytes=4096)
                  for r in range(2)])
    for r, c in enumerate(comms):
        assert c.generation == 0 and c.handles == [b"h0-1048576", b"h1-1048576"]
        assert c._rdv[1] == f"{scope}/g0" and c._rdv_gen == (scope, 0) and c.nvls_bytes == 4096
        assert c.nvls_calls == [(f"{scope}/g0", scope, 0)]
    _run([c._exit_barrier for c in comms])             # the exit barrier of the generation works
    # rank 1 dies and comes back while rank 0 (restarted too) is already exchanging handles in
    # generation 1 with ... nobody: it is alone there until the second restart of rank 1 opens
    # generation 2; its wait ends with Superseded inside rendezvous_comm and it follows
    n_before = len(_ABANDONED)
    res = {}
    t0 = threading.Thread(target=lambda: res.setdefault(0, rendezvous_comm(
        srv.addr, srv.uds_path, scope, 0, 2, 1 << 20, nvls_bytes=4096)))
    t0.start()                                          # new identity for rank 0 -> generation 1
    time.sleep(0.3)
    first = RendezvousClient(srv.addr)                  # rank 1's first restart joins generation 1 ...
    assert first.join(scope, 1, 2, timeout=10) == 1
    time.sleep(0.3)                                     # ... rank 0 is now past join, waiting for h1
    first.close()                                       # ... and dies before publishing its handle
    t1 = threading.Thread(target=lambda: res.setdefault(1, rendezvous_comm(
        srv.addr, srv.uds_path, scope, 1, 2, 1 << 20, nvls_bytes=4096)))
    t1.start()                                          # second restart of rank 1 -> generation 2
    t0.join(20)
    t1.join(20)
    assert res[0].generation == 2 and res[1].generation == 2
    assert res[0].handles == res[1].handles == [b"h0-1048576", b"h1-1048576"]
    assert res[0].nvls_bytes == 4096 and res[0]._rdv_gen == (scope, 2)
    assert len(_ABANDONED) == n_before                  # create() itself was interrupted: nothing to keep
