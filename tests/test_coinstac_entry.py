"""R1: ``entry.py`` → ``coinstac.start(local.run, remote.run)`` (reference ``entry.py:1-5``): both ways the shim can
drive the two callbacks — the JSON-lines server a container adapter talks to, and the one-process simulation."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENTRY = os.path.join(ROOT, "computation", "entry.py")


def _env(**kw):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), DINUNET_QUIET="1",
               CUDA_VISIBLE_DEVICES="")
    env.update(kw)
    return env


def test_json_lines_server_answers_and_survives_bad_input(tmp_path):
    from dinunet_implementations_b200.data import synthetic
    root = str(tmp_path / "fs")
    spec = synthetic.make_fs_dataset(root, site_sizes=(30,), seed=0)
    out = str(tmp_path / "out")
    os.makedirs(out)
    state = {"baseDirectory": os.path.join(root, "input/local0/simulatorRun"), "outputDirectory": out,
             "cacheDirectory": out, "transferDirectory": out, "clientId": "local0", "iteration": 1}
    first = {"mode": "local", "data": {"input": dict(spec[0], epochs=1, num_workers=0), "state": state}}
    lines = json.dumps(first) + "\n" + "this is not json\n" + json.dumps({"mode": "remote", "data": {}}) + "\n"
    r = subprocess.run([sys.executable, ENTRY], input=lines, capture_output=True, text=True, timeout=300, env=_env())
    assert r.returncode == 0, r.stderr[-2000:]
    answers = [json.loads(l) for l in r.stdout.splitlines() if l.strip().startswith("{")]
    assert len(answers) >= 2
    a0 = answers[0]
    assert "output" in a0 and a0["output"].get("phase") == "init_runs", a0           # the site's first report
    assert a0["output"].get("data_size") == 30 or "data_size" in json.dumps(a0)
    assert any("error" in a for a in answers[1:]), "a bad request is answered with an error line, the server goes on"


def test_simulate_mode_runs_to_success(tmp_path):
    from dinunet_implementations_b200.data import synthetic
    root = str(tmp_path / "fs")
    synthetic.make_fs_dataset(root, site_sizes=(40,), seed=0, extra_inputs={"epochs": 2, "num_workers": 0, "num_reducers": 0})
    work = str(tmp_path / "work")
    r = subprocess.run([sys.executable, ENTRY], capture_output=True, text=True, timeout=600,
                       env=_env(DINUNET_SIMULATE=root, DINUNET_SITES="1", DINUNET_WORK_DIR=work))
    assert r.returncode == 0, r.stderr[-2000:]
    log = os.path.join(work, "output/remote/simulatorRun/FS-Classification/fold_0/logs.json")
    assert os.path.exists(log)
    assert len(json.load(open(log))["test_metrics"][0]) == 2
