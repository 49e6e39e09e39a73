"""The supervising launcher (dps_b200.launch.supervise): restart policy of the on-box job, exercised without GPUs through
a fake `run_once` (reference: the "automatic failover" the README claims, README.md:51, which ECS task restarts with a
fresh out-of-range worker id never delivered)."""
from dps_b200 import launch as LN
from dps_b200.config import parse_server_cli


def test_flag_helpers():
    args = ["--mode", "sync", "--inject-fault", "1:12", "--checkpoint=ck.pt", "--resume", "old.pt", "--epochs", "3"]
    assert LN._flag_value(args, "--checkpoint") == "ck.pt"
    assert LN._flag_value(args, "--resume") == "old.pt"
    assert LN._flag_value(args, "--metrics-dir") is None
    out = LN._strip_flag(LN._strip_flag(args, "--inject-fault"), "--resume")
    assert out == ["--mode", "sync", "--checkpoint=ck.pt", "--epochs", "3"]


def test_supervise_restarts_from_the_checkpoint_and_injects_the_fault_once(tmp_path, capsys):
    ck = tmp_path / "job.pt"
    calls = []

    def run_once(args, attempt):
        calls.append((list(args), attempt))
        if attempt == 0:
            ck.write_bytes(b"x")           # the first incarnation got as far as a periodic checkpoint, then a rank died
            return 17
        return 0

    rc = LN.supervise(run_once, ["--mode", "sync", "--checkpoint", str(ck), "--inject-fault", "1:12"], max_restarts=2)
    assert rc == 0 and len(calls) == 2
    first, second = calls[0][0], calls[1][0]
    assert "--inject-fault" in first and "--resume" not in first
    assert "--inject-fault" not in second                                  # the injected fault happens once
    assert second[-2:] == ["--resume", str(ck)] and calls[1][1] == 1      # same arguments + resume, next rendezvous port
    out = capsys.readouterr().out
    assert "RECOVERY attempt 1" in out and "failure_detected_at=" in out


def test_supervise_without_a_checkpoint_restarts_from_scratch_and_gives_up_after_the_budget(tmp_path, capsys):
    calls = []

    def run_once(args, attempt):
        calls.append(list(args))
        return 18

    rc = LN.supervise(run_once, ["--mode", "sync", "--checkpoint", str(tmp_path / "never_written.pt")], max_restarts=2)
    assert rc == 18 and len(calls) == 3                                    # first run + two restarts
    assert all("--resume" not in c for c in calls)
    assert "restart budget is spent" in capsys.readouterr().out


def test_round2_flags_reach_the_config():
    c = parse_server_cli(["--serve-port", "0", "--remote-slots", "2", "--staleness-policy", "world",
                          "--heartbeat-timeout-s", "5", "--no-watchdog", "--checkpoint-every", "7"])
    assert (c.serve_port, c.remote_slots, c.staleness_policy) == (0, 2, "world")
    assert c.heartbeat_timeout_s == 5.0 and c.watchdog is False and c.checkpoint_every == 7
