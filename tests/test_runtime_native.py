"""Native TCP rendezvous store (csrc/runtime/store.cpp, the PMIx-modex analogue): set / blocking get / barrier / all_gather
across processes, on CPU."""
import json
import os
import socket
import time

import pytest
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _proc(rank, world, port, out_dir):
    from ewdml.runtime import lib as L
    if rank == 2:
        time.sleep(0.3)                           # a late joiner: the others block in get / barrier until it arrives
    st = L.Store("127.0.0.1", port, rank, world, timeout_s=30.0)
    st.set(f"key/{rank}", bytes([rank]) * (100 + rank))
    got = {r: st.get(f"key/{r}") for r in range(world)}          # blocks until every key exists
    st.barrier("phase1")
    handles = st.all_gather("handles", f"ipc-handle-of-{rank}".encode().ljust(64, b"\0"))
    big = os.urandom(40000) if rank == 0 else b""
    if rank == 0:
        st.set("big", big)
    blob = st.get("big", cap=1 << 17)
    st.barrier("phase2")
    json.dump({"lens": {str(r): len(v) for r, v in got.items()}, "first": {str(r): v[0] for r, v in got.items()},
               "handles": [h.rstrip(b"\0").decode() for h in handles], "big": len(blob)},
              open(os.path.join(out_dir, f"r{rank}.json"), "w"))
    st.close()


def test_store_set_get_barrier_all_gather(tmp_path):
    world = 3
    mp.spawn(_proc, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = json.load(open(tmp_path / f"r{r}.json"))
        assert res["lens"] == {"0": 100, "1": 101, "2": 102} and res["first"] == {"0": 0, "1": 1, "2": 2}
        assert res["handles"] == [f"ipc-handle-of-{i}" for i in range(world)]
        assert res["big"] == 40000


def test_counters_and_peer_matrix():
    import ctypes as C
    from ewdml.runtime import lib as L
    lib = L.load()
    lib.ewdml_counters_reset()
    lib.ewdml_counter_add(L.CTR_BYTES_SENT, 1234)
    lib.ewdml_counter_add(L.CTR_SEND_CALLS, 2)
    lib.ewdml_peer_matrix_add(3, 1000); lib.ewdml_peer_matrix_add(3, 24); lib.ewdml_peer_matrix_add(0, 7)
    row = (C.c_int64 * L.MAX_PEERS)()
    L.check(lib.ewdml_peer_matrix_get(row))
    assert lib.ewdml_counter_get(L.CTR_BYTES_SENT) == 1234 and lib.ewdml_counter_get(L.CTR_SEND_CALLS) == 2
    assert row[3] == 1024 and row[0] == 7 and row[1] == 0
    lib.ewdml_counters_reset()
    assert lib.ewdml_counter_get(L.CTR_BYTES_SENT) == 0


def _engine(world, rank=0):
    import ctypes as C
    from ewdml.runtime import lib as L
    lib = L.load()
    fake = (C.c_void_p * world)(*[0x1000 * (i + 1) for i in range(world)])     # never dereferenced: no launch happens here
    h = C.c_void_p()
    L.check(lib.ewdml_engine_create(C.byref(h), rank, world, fake, fake, 1 << 20))
    return lib, h


def test_collective_engine_decision_table():
    """coll/tuned analogue: allreduce is one-shot at W <= 2 or at most 256 KiB, two-shot above; forced algorithm wins;
    every other operation has one algorithm."""
    from ewdml.runtime import lib as L
    lib, h2 = _engine(2)
    lib, h8 = _engine(8)
    sel = lambda h, op, n: lib.ewdml_engine_select_algo(h, op, n)
    assert sel(h2, L.COLL_ALLREDUCE, 1 << 30) == L.ALGO_ONESHOT
    assert sel(h8, L.COLL_ALLREDUCE, 1 << 18) == L.ALGO_ONESHOT and sel(h8, L.COLL_ALLREDUCE, (1 << 18) + 16) == L.ALGO_TWOSHOT
    assert sel(h8, L.COLL_ALLGATHER, 1 << 30) == L.ALGO_ONESHOT and sel(h8, L.COLL_BCAST, 16) == L.ALGO_ONESHOT
    L.check(lib.ewdml_engine_set_algo(h8, L.COLL_ALLREDUCE, L.ALGO_ONESHOT))
    assert sel(h8, L.COLL_ALLREDUCE, 1 << 30) == L.ALGO_ONESHOT
    assert lib.ewdml_engine_set_algo(h8, 99, 0) != 0 and lib.ewdml_engine_slot_payload(h8) == 1 << 20
    assert lib.ewdml_engine_window_bytes(8, 1 << 20) == 2 * 2 * 8 * ((1 << 20) + 256)
    assert lib.ewdml_coll_op_name(L.COLL_REDUCE_SCATTER) == b"reduce_scatter"
    lib.ewdml_engine_destroy(h2); lib.ewdml_engine_destroy(h8)


def test_collective_engine_environment_override():
    import subprocess, sys
    code = ("import sys, ctypes as C; sys.path.insert(0, %r); from ewdml.runtime import lib as L; lib = L.load(); "
            "fake = (C.c_void_p * 8)(*[4096 * (i + 1) for i in range(8)]); h = C.c_void_p(); "
            "L.check(lib.ewdml_engine_create(C.byref(h), 0, 8, fake, fake, 1 << 20)); "
            "print(lib.ewdml_engine_select_algo(h, L.COLL_ALLREDUCE, 64), lib.ewdml_engine_select_algo(h, L.COLL_ALLREDUCE, 1 << 30))")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    run = lambda env: subprocess.run([sys.executable, "-c", code % root], capture_output=True, text=True, timeout=120, cwd=root,
                                     env=dict(os.environ, **env)).stdout.split()
    assert run({"EWDML_COLL_ALLREDUCE_ALGO": "twoshot"}) == ["2", "2"]
    assert run({"EWDML_COLL_ONESHOT_MAX_BYTES": "32"}) == ["2", "2"]
    assert run({"EWDML_COLL_ONESHOT_MAX_BYTES": str(1 << 31)}) == ["1", "1"]


def _fdx_proc(rank, world, job, out_dir):
    """Every rank sends a pipe's write end to every other rank, tagged; receives in the OPPOSITE order it was sent in (the
    stash has to hold the early ones); then writes its rank through every descriptor it received."""
    import ctypes as C
    from ewdml.runtime import lib as L
    lib = L.load()
    h = C.c_void_p()
    L.check(lib.ewdml_fdx_open(C.byref(h), job.encode(), rank, world), "fdx_open")
    time.sleep(0.2 * rank)                                   # staggered start: senders meet receivers that are not bound yet
    r_end, w_end = os.pipe()
    for dst in range(world):
        if dst != rank:
            for tag in (7, 8):
                L.check(lib.ewdml_fdx_send(h, dst, w_end, (tag << 8) | 0, 30.0), "fdx_send")
    got = {}
    for tag in (8, 7):                                       # opposite order
        for src in reversed(range(world)):
            if src != rank:
                fd, who = C.c_int(-1), C.c_int(-1)
                L.check(lib.ewdml_fdx_recv(h, (tag << 8) | 0, src, C.byref(fd), C.byref(who), 30.0), "fdx_recv")
                assert who.value == src and fd.value >= 0
                got[(tag, src)] = fd.value
    for (tag, src), fd in got.items():
        os.write(fd, bytes([rank, tag]))                     # lands in rank `src`'s pipe: the descriptor really is its pipe
        os.close(fd)
    os.close(w_end)
    data = b""
    while len(data) < 2 * 2 * (world - 1):
        data += os.read(r_end, 64)
    pairs = sorted((data[i], data[i + 1]) for i in range(0, len(data), 2))
    json.dump({"pairs": pairs}, open(os.path.join(out_dir, f"fdx{rank}.json"), "w"))
    # a receive that nobody answers times out instead of hanging
    fd = C.c_int(-1)
    rc = lib.ewdml_fdx_recv(h, 999, -1, C.byref(fd), None, 0.3)
    assert rc == -3, rc                                      # EWDML_ERR_TIMEOUT
    lib.ewdml_fdx_close(h)


def test_fd_channel_passes_descriptors_between_processes(tmp_path):
    """csrc/runtime/vmm.cu's descriptor channel (what carries the cuMemCreate / multicast-object handles between the ranks of a
    node): SCM_RIGHTS over abstract UNIX datagram sockets, tag + source matching, out-of-order receives, time-outs.  No GPU."""
    import uuid
    world = 3
    mp.spawn(_fdx_proc, args=(world, uuid.uuid4().hex[:12], str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = json.load(open(tmp_path / f"fdx{r}.json"))
        others = [o for o in range(world) if o != r]
        assert res["pairs"] == sorted([o, t] for o in others for t in (7, 8))


def _store_edge(rank, world, port, out_dir):
    os.environ["EWDML_STORE_TIMEOUT_S"] = "1.5"
    from ewdml.runtime import lib as L
    st = L.Store("127.0.0.1", port, rank, world, timeout_s=30.0)
    res = {}
    if rank == 0:
        st.set("large", os.urandom(3 << 20))                 # larger than the default first-guess buffer of get()
    res["large"] = len(st.get("large"))
    st.barrier("a")
    if rank == 1:
        t0 = time.time()
        try:
            st.get("never-set")                              # nobody sets it: the deadline turns the wait into an error
            res["timeout"] = "returned"
        except L.NativeError as e:
            res["timeout"] = "timed out" if "timed out" in str(e) else str(e)
        res["waited"] = time.time() - t0
    json.dump(res, open(os.path.join(out_dir, f"edge{rank}.json"), "w"))
    if rank != 1:
        st.close()


def test_store_large_values_and_deadline(tmp_path):
    """A value beyond get()'s first-guess buffer is fetched again at its real size; a blocking get whose key never appears fails
    with a time-out after EWDML_STORE_TIMEOUT_S instead of hanging the rank (a dead peer must not hang the job)."""
    world = 2
    mp.spawn(_store_edge, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (json.load(open(tmp_path / f"edge{r}.json")) for r in (0, 1))
    assert r0["large"] == r1["large"] == 3 << 20
    assert r1["timeout"] == "timed out" and 1.0 < r1["waited"] < 10.0


@pytest.mark.parametrize("sanitizer", ["thread", "address,undefined"])
def test_store_under_sanitizers(tmp_path, sanitizer):
    """The store's server and client code under ThreadSanitizer and Address/UB sanitizer: several ranks as threads, values
    that cross the buffer-resize path, named barriers, and the host leaving first (tests/native/store_stress.cpp)."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # Can this machine run a sanitized binary at all?  (No runtime library installed; a kernel whose address-space layout
    # the sanitizer runtime of this compiler does not know: "unexpected memory mapping".)  If not, that is not a finding.
    probe_src, probe = tmp_path / "probe.cpp", str(tmp_path / "probe")
    probe_src.write_text("#include <thread>\nint main(){int x=0; std::thread t([&]{x=1;}); t.join(); return x-1;}\n")
    pb = subprocess.run(["g++", "-std=c++17", f"-fsanitize={sanitizer}", str(probe_src), "-lpthread", "-o", probe],
                        capture_output=True, text=True)
    if pb.returncode != 0:
        pytest.skip("cannot build with -fsanitize=" + sanitizer)
    try:
        pr = subprocess.run([probe], capture_output=True, text=True, timeout=60)
    except subprocess.TimeoutExpired:
        pytest.skip("sanitized probe does not run here")
    if pr.returncode != 0:
        pytest.skip("sanitized probe does not run here: " + (pr.stderr.strip().splitlines() or ["?"])[0][:120])
    exe = str(tmp_path / "store_stress")
    build = subprocess.run(["g++", "-std=c++17", "-O1", "-g", f"-fsanitize={sanitizer}", "-fno-omit-frame-pointer",
                            os.path.join(root, "tests", "native", "store_stress.cpp"),
                            os.path.join(root, "csrc", "runtime", "store.cpp"), "-lpthread", "-o", exe],
                           capture_output=True, text=True)
    if build.returncode != 0 and "sanitizer" in build.stderr.lower() and "cannot find" in build.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert build.returncode == 0, build.stderr[-2000:]
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1 exitcode=66", ASAN_OPTIONS="detect_leaks=1 exitcode=67",
               UBSAN_OPTIONS="halt_on_error=1 print_stacktrace=1")
    env.pop("LD_PRELOAD", None)
    run = subprocess.run([exe, str(_free_port()), "6", "40"], capture_output=True, text=True, timeout=300, env=env)
    assert run.returncode == 0, (run.stdout + run.stderr)[-3000:]
    assert "store_stress ok" in run.stdout
    assert "Sanitizer" not in run.stderr, run.stderr[-3000:]


def test_c_abi_header_library_and_binding_agree():
    """Every function declared in include/ewdml.h is exported by libewdml.so, everything the library exports beyond that is
    one of the two internal helpers, and every function the ctypes binding names exists (the C ABI is the public surface
    for non-Python users)."""
    import re
    import shutil
    import subprocess
    from ewdml.runtime import lib as L
    if shutil.which("nm") is None:
        pytest.skip("no nm")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    L.load()
    so = os.path.join(root, "ewdml", "_C", "libewdml.so")
    header = open(os.path.join(root, "include", "ewdml.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    header = re.sub(r"//[^\n]*", "", header)
    declared = set(re.findall(r"\b(ewdml_[a-z0-9_]+)\s*\(", header))
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("ewdml_")}
    assert declared - exported == set(), sorted(declared - exported)
    assert exported - declared <= {"ewdml_device_error_word", "ewdml_spin_limit"}, sorted(exported - declared)
    assert len(declared) > 90
    src = open(os.path.join(root, "ewdml", "runtime", "lib.py")).read()
    bound = set(re.findall(r"\"(ewdml_[a-z0-9_]+)\"\s*:", src))
    assert bound and bound <= exported, sorted(bound - exported)
