"""Flag surface (distributed_nn.py:30-70) and the env-override registry."""
import pytest

from ewdml.config import TrainConfig, parse_args, registry, Registry


def test_reference_flag_surface_and_defaults():
    c = parse_args([])
    assert (c.batch_size, c.test_batch_size, c.epochs, c.max_steps) == (128, 500, 100, 10000)
    assert (c.lr, c.momentum, c.seed, c.log_interval) == (0.01, 0.5, 1, 10)
    assert (c.network, c.dataset, c.mode, c.comm_type) == ("LeNet", "MNIST", "normal", "Bcast")
    assert (c.num_aggregate, c.eval_freq, c.train_dir) == (5, 50, "output/models/")
    assert (c.compress_grad, c.gather_type, c.kill_threshold) == ("compress", "gather", 7.0)


def test_reference_script_command_line_parses():
    # flags of src/run_pytorch_single.sh:4-17, including the empty --enable-gpu=
    c = parse_args(["--lr=0.01", "--momentum=0.9", "--max-steps=10000", "--epochs=10", "--network=LeNet",
                    "--dataset=MNIST", "--batch-size=64", "--comm-type=Bcast", "--num-aggregate=5",
                    "--mode=normal", "--eval-freq=20", "--gather-type=gather", "--compress-grad=compress",
                    "--enable-gpu=", "--train-dir=/tmp/x/", "--local_rank=0"])
    assert c.enable_gpu is False and c.batch_size == 64 and c.local_rank == 0
    assert parse_args(["--enable-gpu=1"]).enable_gpu is True
    assert parse_args(["--enable-gpu"]).enable_gpu is True


def test_methods_resolve():
    assert parse_args(["--method", "1"]).resolved_compressor() == "none"
    assert parse_args(["--method", "1"]).resolved_pull() == "weights"
    m4 = parse_args(["--method", "4"])
    assert m4.resolved_compressor() == "qsgd" and m4.resolved_compress_pull() and m4.resolved_pull() == "grads"
    m6 = parse_args(["--method", "6"])
    assert m6.resolved_compressor() == "topk-qsgd" and m6.resolved_sync_period() == 20
    assert parse_args(["--compress-grad", "none"]).resolved_compressor() == "none"
    with pytest.raises(ValueError):
        TrainConfig(network="AlexNet").validate()


def test_env_override_precedence(monkeypatch):
    monkeypatch.setenv("EWDML_QSGD_LEVELS", "64")
    assert parse_args([]).qsgd_levels == 64
    assert parse_args(["--qsgd-levels", "16"]).qsgd_levels == 16      # CLI beats env
    monkeypatch.setenv("EWDML_MY_TUNABLE", "7")
    r = Registry()
    assert r.register("my_tunable", 3, "x") == 7 and r.dump()["my_tunable"]["source"] == "env"
    r.set("my_tunable", "9"); assert r.get("my_tunable") == 9
    assert "spin_limit" in registry.dump()


def test_cluster_tool_on_the_local_pseudo_host(tmp_path):
    """tools/cluster.py (the host-driving half of the reference's pytorch_ec2.py): hosts file, run_command, launch with
    recorded PIDs, status, PID-based kill_all -- exercised on the `local` pseudo host, no ssh involved."""
    import os, subprocess, sys, time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hosts, pids, log = tmp_path / "hosts", tmp_path / "pids", tmp_path / "job.log"
    hosts.write_text("# comment\nlocal 2\n")
    base = [sys.executable, os.path.join(root, "tools", "cluster.py"), "--hosts", str(hosts), "--pids", str(pids)]
    run = lambda *args: subprocess.run(base + list(args), capture_output=True, text=True, timeout=120, cwd=root)
    assert run("get_hosts").stdout.strip() == "local slots=2"
    assert "[local] hello" in run("run_command", "--", "echo", "hello").stdout
    dry = run("--dry-run", "launch", "--", "python", "x.py")
    assert "ewdml.cli.launch -np 2" in dry.stdout and not pids.exists()
    out = run("launch", "--log", str(log), "--master-port", "29761", "--", sys.executable, "-c", "import time; time.sleep(60)")
    assert out.returncode == 0 and pids.exists(), out.stderr
    time.sleep(1.0)
    assert "running" in run("status").stdout
    assert run("kill_all").returncode == 0 and not pids.exists()


def test_symmetric_heap_allocator_is_deterministic_and_coalesces():
    """ewdml_symheap_*: offsets only (no GPU needed); two heaps fed the same sequence agree, frees coalesce."""
    import ctypes as C
    from ewdml.runtime import lib as L
    lib = L.load()

    def run():
        h = C.c_void_p()
        L.check(lib.ewdml_symheap_create(C.byref(h), 1 << 20, 256))
        offs = [lib.ewdml_symheap_alloc(h, n) for n in (1000, 5000, 300)]
        L.check(lib.ewdml_symheap_free(h, offs[1]))
        offs += [lib.ewdml_symheap_alloc(h, 2000), lib.ewdml_symheap_alloc(h, 4000)]
        return h, offs

    (h1, o1), (h2, o2) = run(), run()
    assert o1 == o2 == [0, 1024, 6144, 1024, 6656]               # first fit reuses the hole, the rest goes behind
    assert lib.ewdml_symheap_alloc(h1, 1 << 21) == -1 and b"do not fit" in lib.ewdml_last_error()
    for off in (0, 6144, 1024, 6656):
        L.check(lib.ewdml_symheap_free(h1, off))
    assert lib.ewdml_symheap_used(h1) == 0 and lib.ewdml_symheap_largest_free(h1) == 1 << 20   # fully coalesced again
    assert lib.ewdml_symheap_free(h1, 12345) != 0                 # not a live allocation
    lib.ewdml_symheap_destroy(h1); lib.ewdml_symheap_destroy(h2)


def test_c_abi_header_is_c99_and_links_from_c(tmp_path):
    """include/ewdml.h compiles as pedantic C99 and a plain C program links and runs against libewdml.so (no GPU needed for
    the host-side entry points)."""
    import os, shutil, subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if shutil.which("gcc") is None or not os.path.isfile(os.path.join(root, "ewdml", "_C", "libewdml.so")):
        import pytest
        pytest.skip("needs gcc and the built library")
    exe = str(tmp_path / "c_abi_demo")
    cc = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(root, "include"),
                         os.path.join(root, "examples", "c_abi_demo.c"), "-o", exe, "-L" + os.path.join(root, "ewdml", "_C"), "-lewdml",
                         "-Wl,-rpath," + os.path.join(root, "ewdml", "_C")], capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr
    out = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and "offsets 0 1024 used 71168" in out.stdout and "counter 3" in out.stdout


def test_cloud_tool_plan_and_dry_run(tmp_path):
    """tools/cloud.py (the instance-renting half of the reference's pytorch_ec2.py): self-referential configuration, the
    run_instances requests for master + workers, dry-run of launch / hosts / terminate, hosts-file format -- all offline."""
    import importlib.util, json, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    tool = os.path.join(root, "tools", "cloud.py")
    cfgf = tmp_path / "c.json"
    cfgf.write_text(json.dumps({"name": "t1", "region": "eu-west-1", "image_id": "ami-x", "key_name": "k", "n_workers": 3,
                                "master_type": "p6-b200.48xlarge", "spot_price": "9.5", "hosts_file": str(tmp_path / "hosts")}))
    out = subprocess.run([sys.executable, tool, "--config", str(cfgf), "plan"], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stderr
    plan = json.loads(out.stdout)
    assert plan["config"]["availability_zone"] == "eu-west-1b" and plan["config"]["worker_type"] == "p6-b200.48xlarge"
    reqs = plan["run_instances"]
    assert [r["MaxCount"] for r in reqs] == [1, 3]
    assert reqs[1]["InstanceMarketOptions"]["SpotOptions"]["MaxPrice"] == "9.5"
    assert {t["Value"] for t in reqs[0]["TagSpecifications"][0]["Tags"]} == {"t1", "master"}
    dry = subprocess.run([sys.executable, tool, "--config", str(cfgf), "--dry-run", "launch"], capture_output=True, text=True, timeout=60)
    assert dry.returncode == 0 and dry.stdout.count("ec2.run_instances(") == 2
    assert subprocess.run([sys.executable, tool, "--config", str(cfgf), "--dry-run", "terminate"], capture_output=True, text=True,
                          timeout=60).returncode == 0
    # hosts file in the format tools/cluster.py reads (master first)
    spec = importlib.util.spec_from_file_location("ewdml_cloud_tool", tool)
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    cfg = mod.load_config(str(cfgf))
    mod.write_hosts([{"id": "i-1", "role": "master", "private_ip": "10.0.0.1", "public_ip": ""},
                     {"id": "i-2", "role": "worker", "private_ip": "10.0.0.2", "public_ip": ""}], cfg)
    spec2 = importlib.util.spec_from_file_location("ewdml_cluster_tool", os.path.join(root, "tools", "cluster.py"))
    cl = importlib.util.module_from_spec(spec2); spec2.loader.exec_module(cl)
    assert cl.read_hosts(str(tmp_path / "hosts")) == [("ubuntu@10.0.0.1", 8), ("ubuntu@10.0.0.2", 8)]
    import pytest
    with pytest.raises(ValueError):
        mod.resolve({"a": "%(b)s", "b": "%(a)sx"})


def test_validate_rejects_contradictory_or_out_of_range_flags():
    import pytest
    from ewdml.config import TrainConfig
    TrainConfig().validate()
    for kw in (dict(overlap=True, error_feedback=True), dict(batch_size=0), dict(momentum=1.0), dict(lr=-0.1),
               dict(topk_ratio=0.0), dict(qsgd_levels=0), dict(mode="sometimes"), dict(network="AlexNet")):
        with pytest.raises(ValueError):
            TrainConfig(**kw).validate()


def test_report_tool_reduces_metrics_files_to_the_reports_table(tmp_path, capsys):
    """ewdml.cli.report: accuracy / communication vs compute minutes / MB per iteration / steps to a target accuracy
    (Report.zip!main.tex:143-158) from --metrics-file records; the slowest worker defines the times."""
    import json
    from ewdml.cli import report

    def write(path, mb, comm_ms, acc_at):
        with open(path, "w") as f:
            for step in range(20):
                for rank, slow in ((1, 1.0), (2, 2.0)):
                    f.write(json.dumps({"rank": rank, "step": step, "loss": 2.0 / (step + 1), "prec1": acc_at(step), "prec5": 100.0,
                                        "bytes_sent": int(mb * 5e5), "bytes_recv": int(mb * 5e5), "compression_ratio": 4.0 / mb,
                                        "ms_send": comm_ms * slow / 2, "ms_recv": comm_ms * slow / 2, "ms_compute": 60.0 * slow,
                                        "time": 1000.0 + step * 0.2}) + "\n")
            f.write('{"rank": 1, "step": 20, "lo')            # a record cut short by a killed job is skipped

    dense, sparse = str(tmp_path / "dense.jsonl"), str(tmp_path / "sparse.jsonl")
    write(dense, 4.0, 30.0, lambda s: min(100.0, 10.0 * s))
    write(sparse, 0.5, 6.0, lambda s: min(100.0, 8.0 * s))
    d = report.summarize(report.load(dense), tail=5, target=90.0)
    s = report.summarize(report.load(sparse), tail=5, target=90.0)
    assert d["workers"] == 2 and d["steps"] == 20 and d["top1_final"] == 100.0
    assert d["mb_per_iteration"] == pytest.approx(4.0) and s["mb_per_iteration"] == pytest.approx(0.5)
    assert d["comm_minutes"] == pytest.approx(20 * 60.0 / 60e3) and d["compute_minutes"] == pytest.approx(20 * 120.0 / 60e3)   # rank 2
    assert d["comm_fraction"] == pytest.approx(60.0 / 180.0)
    assert d["wall_minutes"] == pytest.approx(19 * 0.2 / 60.0)
    assert d["steps_to_target"] == 12 and s["steps_to_target"] == 14        # first window of 5 whose mean reaches 90 %
    assert report.main([f"dense={dense}", f"sparse={sparse}", "--tail", "5", "--target", "90"]) == 0
    out = capsys.readouterr().out
    assert "dense" in out and "sparse" in out and "8.0x" in out and "steps to 90 %" in out
    assert report.main([dense, "--json"]) == 0
    assert json.loads(capsys.readouterr().out)[0]["run"] == dense
    assert report.summarize([]) == {"steps": 0, "workers": 0}
