"""W15/W20: aggregate schema, phantom-worker hygiene, headless PNG charts -- also fed with the reference's own
published result files (which its own visualiser cannot plot because server_metrics is null)."""
import glob
import json
import os

from dps_b200.utils import charts as CH
from dps_b200.utils import metrics as M
from dps_b200.utils.png import Canvas, read_png_size


def _worker(wid, steps=10, acc=(1.0, 2.0), t=(5.0, 6.0), ips=100.0):
    return M.worker_final_metrics(worker_id=wid, total_workers=2, total_time=sum(t) + 1, epoch_times=list(t),
                                  accuracies=list(acc), local_steps=steps, batch_size=128, learning_rate=0.1,
                                  num_epochs=len(t), extra={"images_per_second": ips})


def test_aggregate_excludes_phantom_and_duplicate_workers():
    recs = [_worker(0), _worker(1), _worker(1, acc=(3.0, 4.0)), _worker(9, steps=0, acc=(), t=())]
    agg = M.aggregate_workers(recs)
    assert agg["num_workers"] == 2 and agg["total_local_steps"] == 20
    assert agg["final_test_accuracy_percent"] == (2.0 + 4.0) / 2
    assert agg["epoch_times_by_epoch"][1] == {"epoch": 2, "max_time": 6.0, "avg_time": 6.0, "min_time": 6.0}
    assert agg["images_per_second_sum"] == 200.0


def test_parse_metrics_lines_contract():
    lines = ["noise", M.METRICS_PREFIX + " " + json.dumps(_worker(0)),
             "x METRICS_JSON: " + json.dumps({"type": "SERVER_FINAL_METRICS", "mode": "sync"}), "METRICS_JSON: {bad"]
    out = M.parse_metrics_lines(lines)
    assert len(out["worker"]) == 1 and len(out["server"]) == 1


def test_png_writer(tmp_path):
    cv = Canvas(64, 32)
    cv.line(0, 0, 63, 31, (255, 0, 0), 2)
    cv.text(2, 2, "AB 12")
    p = str(tmp_path / "a.png")
    cv.save(p)
    assert read_png_size(p) == (64, 32) and os.path.getsize(p) > 60


def test_collect_and_plot_cli(tmp_path):
    for mode in ("sync", "async"):
        for w in (2, 4):
            d = tmp_path / f"{mode}{w}"
            d.mkdir()
            for i in range(w):
                M.emit(_worker(i, ips=100.0 * (1 + 0.1 * i)), str(d), "worker")
            M.emit(M.server_final_metrics(mode=mode, total_workers=w, total_time=12, global_step=10, total_updates=10,
                                          gradients_processed=10 * w, avg_update_time=0.01, learning_rate=0.1), str(d),
                   "server")
            assert CH.main(["collect", "--metrics-dir", str(d), "--name", f"{mode}_{w}workers", "--output",
                            str(tmp_path / "results" / f"{mode}_{w}workers.json")]) == 0
    assert CH.main(["plot", "--results-dir", str(tmp_path / "results"), "--output-dir", str(tmp_path / "charts")]) == 0
    pngs = sorted(os.listdir(tmp_path / "charts"))
    assert pngs == ["scaling.png", "sync_vs_async_2workers.png", "sync_vs_async_4workers.png"]
    assert read_png_size(str(tmp_path / "charts" / "scaling.png")) == (1000, 760)


def test_plots_reference_published_results(tmp_path):
    files = sorted(glob.glob("/root/reference/experiment_results/*.json"))
    if not files:
        import pytest
        pytest.skip("reference results not mounted")
    results = CH.load_results(files=files)
    assert len(results) == 3                      # sync_4, async_4, async_8
    keys = sorted(CH.result_key(r)[0] for r in results)
    assert keys == ["async", "async", "sync"]
    CH.plot_scaling(results, str(tmp_path / "ref_scaling.png"))
    assert read_png_size(str(tmp_path / "ref_scaling.png")) == (1000, 760)
