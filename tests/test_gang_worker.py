"""Gang jobs through the job API (SURVEY.md sections 2.3, 5.3, 5.8): the hive client leases
1 / 2 / 4 / 8 ranks per job from the RankGroup, the ranks run the job SPMD, the gang's first rank answers.
CPU tier: two real rank processes over gloo with the eager back end; the GPU tier repeats it on NCCL with the
sm_100a patch engine (tests/test_dist_gpu.py)."""
import asyncio
import multiprocessing as mp
import os
import time

import numpy as np
import pytest

from tests.helpers import decode_artifact_image
from tests.test_worker_hive import FakeHive


def _settings(**kw):
    from chiaswarm_b200.settings import Settings

    s = Settings()
    s.sdaas_uri, s.sdaas_token, s.worker_name = "http://127.0.0.1:9", "secret", "gangrig"
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def test_plan_gang_size_policy():
    from chiaswarm_b200.parallel.gang_worker import aligned_subsets, plan_gang_size

    dp = _settings(parallel_mode="dp", gang_size=8)
    gang = _settings(parallel_mode="gang", gang_size=8)
    sdxl = {"model_name": "stabilityai/stable-diffusion-xl-base-1.0", "height": 1024, "width": 1024}
    assert plan_gang_size(sdxl, dp, 8) == 1                    # reference behaviour: one GPU per job
    assert plan_gang_size(sdxl, gang, 8) == 8                  # one large image: patch-parallel over the gang
    assert plan_gang_size(sdxl, gang, 4) == 4
    assert plan_gang_size(dict(sdxl, height=512, width=512), gang, 8) == 1   # small image: launch-bound, stays DP
    assert plan_gang_size(dict(sdxl, num_images_per_prompt=8, height=768, width=768), gang, 8) == 8
    assert plan_gang_size(dict(sdxl, num_images_per_prompt=6), gang, 8) == 2  # largest power of two dividing 6
    assert plan_gang_size(dict(sdxl, num_images_per_prompt=3), gang, 8) == 1
    assert plan_gang_size({"workflow": "vid2vid", "model_name": "timbrooks/instruct-pix2pix"}, gang, 8) == 8
    assert plan_gang_size({"workflow": "txt2vid", "model_name": "m"}, _settings(parallel_mode="gang", gang_size=4), 8) == 4
    assert plan_gang_size({"workflow": "stitch", "model_name": "m"}, gang, 8) == 1
    assert plan_gang_size(dict(sdxl, start_image_uri="x"), gang, 8) == 1
    assert plan_gang_size(sdxl, _settings(parallel_mode="gang", gang_size=3), 8) == 2  # rounded down to a buddy size
    assert aligned_subsets(4) == [(0, 1), (2, 3), (0, 1, 2, 3)]
    assert len(aligned_subsets(8)) == 4 + 2 + 1


def _job(job_id, **kw):
    job = {"id": job_id, "model_name": "tiny-sd", "prompt": "a gang job", "height": 64, "width": 64,
           "num_inference_steps": 2, "seed": 11, "content_type": "image/png"}
    job.update(kw)
    return job


@pytest.fixture
def gang_env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    return {"SDAAS_ROOT": str(tmp_path), "SDAAS_RANDOM_INIT": "1", "SDAAS_KERNEL_BACKEND": "torch"}


def test_gang_jobs_match_single_rank(gang_env, monkeypatch):
    """A patch-parallel single image and a split batch on a 2-rank gang give the images one rank gives."""
    from chiaswarm_b200.generator import synchronous_do_work_function
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    for k, v in gang_env.items():
        monkeypatch.setenv(k, v)
    single = _job("one")
    batch = _job("four", num_images_per_prompt=4, seed=5)
    xl = _job("xl", model_name="tiny-sdxl", seed=11)    # the SDXL path: pooled text + size conditioning under a row split
    ref_single = synchronous_do_work_function(dict(single), Device("cpu"))
    ref_batch = synchronous_do_work_function(dict(batch), Device("cpu"))
    ref_xl = synchronous_do_work_function(dict(xl), Device("cpu"))
    assert "error" not in ref_single["pipeline_config"], ref_single["pipeline_config"]

    sched = GangScheduler(2, _settings(parallel_mode="gang", gang_size=2, patch_min_pixels=64 * 64), use_cuda=False,
                          env=gang_env, job_timeout_s=300)
    sched.start()
    try:
        got_single = sched.submit(single)
        got_batch = sched.submit(batch)
        got_xl = sched.submit(xl)
        assert sched.leases_by_size == {2: 3}
        assert sched.pool.free_count() == 2
    finally:
        sched.stop()
    for ref, got in ((ref_single, got_single), (ref_batch, got_batch), (ref_xl, got_xl)):
        assert "error" not in got["pipeline_config"], got["pipeline_config"]
        assert got["pipeline_config"]["seed"] == ref["pipeline_config"]["seed"]
        a = np.asarray(decode_artifact_image(ref["artifacts"]["primary"]), dtype=np.float32)
        b = np.asarray(decode_artifact_image(got["artifacts"]["primary"]), dtype=np.float32)
        assert a.shape == b.shape
        assert np.abs(a - b).max() <= 2.0  # same latents up to fp32 reduction order -> same 8-bit image +- 1-2 levels


def test_gang_worker_serves_a_fake_hive(gang_env):
    from chiaswarm_b200.parallel.gang_worker import GangScheduler, GangWorker

    async def main():
        hive = FakeHive()
        uri = await hive.start()
        hive.jobs = [_job("patch"), _job("dp", height=32, width=32), {"id": "bad", "model_name": "tiny-sd", "height": 5000, "width": 64}]
        s = _settings(parallel_mode="gang", gang_size=2, patch_min_pixels=64 * 64, sdaas_uri=uri)
        sched = GangScheduler(2, s, use_cuda=False, env=gang_env, job_timeout_s=300)
        try:
            w = GangWorker(s, scheduler=sched, cpu_lane=False)

            async def fast_sleep(_):
                await asyncio.sleep(0)

            await w.run(max_polls=2, sleep=fast_sleep)
            by_id = {r["id"]: r for r in hive.results}
            assert sorted(by_id) == ["bad", "dp", "patch"]
            assert by_id["bad"]["fatal_error"] is True
            assert "error" not in by_id["patch"]["pipeline_config"]
            # "patch" and the oversized "bad" job are large enough for a gang of 2, "dp" (32x32) keeps one rank
            assert sched.leases_by_size.get(2) == 2 and sched.leases_by_size.get(1) == 1
        finally:
            sched.stop()
            await hive.stop()

    asyncio.run(main())


def test_vid2vid_frames_scatter_over_a_gang(gang_env, tmp_path, monkeypatch):
    """Frame-level parallelism: the chunks of a video-to-video job are dealt out over the
    ranks of the leased gang and gathered in order; the video equals the single-rank one frame for frame."""
    import base64

    import cv2

    from chiaswarm_b200.generator import synchronous_do_work_function
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.parallel.gang_worker import GangScheduler
    from tests.test_workflows_cpu import _tiny_video

    env = dict(gang_env, SDAAS_FRAME_BATCH="1", SDAAS_LOCAL_INPUTS="1", SDAAS_LOCAL_INPUT_ROOT=str(tmp_path),
               SDAAS_VIDEO_HEIGHT="96")
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    from chiaswarm_b200.video import pix2pix

    monkeypatch.setattr(pix2pix, "FRAME_BATCH", 1)
    job = {"id": "v2v", "workflow": "vid2vid", "model_name": "tiny-pix2pix", "prompt": "night", "num_inference_steps": 1,
           "video_uri": _tiny_video(tmp_path / "in.mp4", frames=3), "seed": 8}

    def frames_of(result, name):
        assert "error" not in result["pipeline_config"], result["pipeline_config"]
        path = tmp_path / name
        path.write_bytes(base64.b64decode(result["artifacts"]["primary"]["blob"]))
        cap, out = cv2.VideoCapture(str(path)), []
        while True:
            ok, f = cap.read()
            if not ok:
                break
            out.append(f.astype(np.float32))
        return out

    ref = frames_of(synchronous_do_work_function(dict(job), Device("cpu")), "ref.mp4")
    sched = GangScheduler(2, _settings(parallel_mode="gang", gang_size=2), use_cuda=False, env=env, job_timeout_s=600)
    sched.start()
    try:
        got = frames_of(sched.submit(job), "gang.mp4")
        assert sched.leases_by_size == {2: 1}
    finally:
        sched.stop()
    assert len(ref) == len(got) == 3
    # same frames up to fp32 reduction order (the ranks run with fewer threads) seen through a lossy codec
    assert max(float(np.abs(a - b).mean()) for a, b in zip(ref, got)) <= 0.5
    assert max(float(np.abs(a - b).max()) for a, b in zip(ref, got)) <= 6.0


def test_txt2vid_on_a_gang_scatters_the_decode(gang_env):
    """Text-to-video on a leased gang (eager back end here: denoising is replicated, the VAE decode of the frames is
    dealt out over the ranks and all-gathered; on GPUs the engine also splits the denoising along the frame axis):
    the clip equals the single-rank one."""
    import base64

    from chiaswarm_b200.generator import synchronous_do_work_function
    from chiaswarm_b200.gpu.device import Device
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    job = {"id": "t2v", "workflow": "txt2vid", "model_name": "tiny-t2v", "prompt": "waves", "num_frames": 5, "height": 32,
           "width": 32, "num_inference_steps": 2, "seed": 9}
    import os

    os.environ.update(gang_env)
    ref = synchronous_do_work_function(dict(job), Device("cpu"))
    sched = GangScheduler(2, _settings(parallel_mode="gang", gang_size=2), use_cuda=False, env=gang_env, job_timeout_s=300)
    sched.start()
    try:
        got = sched.submit(job)
        assert sched.leases_by_size == {2: 1}
    finally:
        sched.stop()
    import cv2

    def frames_of(r, name, tmp=os.environ["SDAAS_ROOT"]):
        assert "error" not in r["pipeline_config"], r["pipeline_config"]
        path = os.path.join(tmp, name)
        with open(path, "wb") as fh:
            fh.write(base64.b64decode(r["artifacts"]["primary"]["blob"]))
        cap, out = cv2.VideoCapture(path), []
        while True:
            ok, f = cap.read()
            if not ok:
                break
            out.append(f.astype(np.float32))
        return out

    a, b = frames_of(ref, "ref.mp4"), frames_of(got, "gang.mp4")
    assert len(a) == len(b) == 5
    # (decode batches of 3 + 2 frames instead of 5: fp rounding, amplified a little by the lossy video codec)
    assert max(float(np.abs(x - y).mean()) for x, y in zip(a, b)) <= 1.0
    assert max(float(np.abs(x - y).max()) for x, y in zip(a, b)) <= 16.0


def test_launcher_picks_the_gang_server(monkeypatch):
    """`settings.parallel_mode` / `gang_size` decide between the per-GPU supervisor and the gang server."""
    from chiaswarm_b200.parallel import gang_worker, launcher

    calls = []
    monkeypatch.setattr(gang_worker, "serve", lambda settings, world: calls.append(("gang", world, settings.gang_size)))
    monkeypatch.setattr(launcher.Supervisor, "run", lambda self, max_seconds=None: calls.append(("dp", len(self.gpus))))
    monkeypatch.setenv("SDAAS_PARALLEL_MODE", "gang")
    monkeypatch.setenv("SDAAS_GANG_SIZE", "4")
    assert launcher.main(["--gpus", "0,1,2,3"]) == 0
    monkeypatch.setenv("SDAAS_PARALLEL_MODE", "dp")
    assert launcher.main(["--gpus", "0,1"]) == 0
    assert calls == [("gang", 4, 4), ("dp", 2)]


def _dying_rank(rank, world, port, job_q, result_q, use_cuda, env=None):
    """Stand-in rank body: rank 1 exits in the middle of its first job (no torch needed)."""
    result_q.put(("ready", rank, None))
    first = not os.path.exists(os.path.join(env["MARK"], f"restarted{rank}"))
    open(os.path.join(env["MARK"], f"restarted{rank}"), "w").close()
    while True:
        msg = job_q.get()
        if msg is None:
            return
        lease_id, ranks, job = msg
        if rank == 1 and first:
            os._exit(7)
        result_q.put(("result" if ranks[0] == rank else "done", lease_id, {"id": job["id"], "ok": True}))


def _stuck_rank(rank, world, port, job_q, result_q, use_cuda, env=None):
    result_q.put(("ready", rank, None))
    while True:
        msg = job_q.get()
        if msg is None:
            return
        time.sleep(3600)


def test_dead_rank_fails_the_job_and_restarts_the_world(tmp_path):
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    s = _settings(parallel_mode="gang", gang_size=2)
    sched = GangScheduler(2, s, use_cuda=False, env={"MARK": str(tmp_path)}, job_timeout_s=60, target=_dying_rank)
    sched.start()
    try:
        r = sched.submit({"id": "j1", "model_name": "m", "workflow": "vid2vid"})
        assert "died" in r["pipeline_config"]["error"] and "fatal_error" not in r  # non-fatal: the hive re-issues it
        assert sched.restarts == 1 and sched.pool.free_count() == 2
        r2 = sched.submit({"id": "j2", "model_name": "m", "workflow": "vid2vid"})  # the new world serves
        assert r2 == {"id": "j2", "ok": True}
    finally:
        sched.stop()


def test_watchdog_times_a_wedged_gang_out(tmp_path):
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    sched = GangScheduler(2, _settings(parallel_mode="gang", gang_size=2), use_cuda=False, env={},
                          job_timeout_s=1.0, target=_stuck_rank)
    sched.start()
    try:
        t0 = time.time()
        r = sched.submit({"id": "j", "model_name": "m", "workflow": "txt2vid"})
        assert "exceeded" in r["pipeline_config"]["error"] and time.time() - t0 < 30
        assert sched.restarts == 1
    finally:
        sched.stop()


def _lone_failure_rank(rank, world, port, job_q, result_q, use_cuda, env=None):
    """Rank 1 comes back with an error of its own; the leader never returns (it waits in a collective for rank 1)."""
    result_q.put(("ready", rank, None))
    first = not os.path.exists(os.path.join(env["MARK"], f"up{rank}"))
    open(os.path.join(env["MARK"], f"up{rank}"), "w").close()
    while True:
        msg = job_q.get()
        if msg is None:
            return
        lease_id, ranks, job = msg
        if first and rank == 1:
            result_q.put(("done", lease_id, {"error": "CUDA out of memory"}))
        elif first:
            time.sleep(3600)
        else:
            result_q.put(("result" if ranks[0] == rank else "done", lease_id, {"id": job["id"], "ok": True}))


def test_a_rank_that_fails_alone_does_not_cost_the_whole_job_timeout(tmp_path):
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    sched = GangScheduler(2, _settings(parallel_mode="gang", gang_size=2), use_cuda=False, env={"MARK": str(tmp_path)},
                          job_timeout_s=600, target=_lone_failure_rank)
    sched.straggler_grace_s = 0.5
    sched.start()
    try:
        t0 = time.time()
        r = sched.submit({"id": "j1", "model_name": "m", "workflow": "txt2vid"})
        assert time.time() - t0 < 60
        err = r["pipeline_config"]["error"]
        assert "rank 1 failed (CUDA out of memory)" in err and "ranks [0] did not return" in err
        assert sched.restarts == 1
        assert sched.submit({"id": "j2", "model_name": "m", "workflow": "txt2vid"}) == {"id": "j2", "ok": True}
    finally:
        sched.stop()


def _env_echo_rank(rank, world, port, job_q, result_q, use_cuda, env=None):
    os.environ.update(env or {})
    result_q.put(("ready", rank, None))
    while True:
        msg = job_q.get()
        if msg is None:
            return
        lease_id, ranks, job = msg
        result_q.put(("result" if ranks[0] == rank else "done", lease_id,
                      {"id": job["id"], "attn": os.environ.get("SW_PATCH_ATTN", "kv")}))


def test_tp_mode_selects_head_parallel_attention_in_the_ranks():
    from chiaswarm_b200.parallel.gang_worker import GangScheduler

    for mode, want in (("tp", "tp"), ("gang", "kv")):
        sched = GangScheduler(2, _settings(parallel_mode=mode, gang_size=2), use_cuda=False, env={}, target=_env_echo_rank)
        sched.start()
        try:
            assert sched.submit({"id": "j", "model_name": "m", "workflow": "txt2vid"})["attn"] == want
        finally:
            sched.stop()


def test_the_reference_start_command_hands_over_to_the_launcher_in_gang_mode(monkeypatch):
    """`python -m swarm.worker` with gangs configured must not quietly serve one job per GPU (reference start command:
    Dockerfile:37)."""
    from chiaswarm_b200 import worker
    from chiaswarm_b200.parallel import launcher
    from chiaswarm_b200.settings import Settings

    s = Settings()
    s.parallel_mode, s.gang_size = "patch", 4
    calls = []
    monkeypatch.setattr(worker, "load_settings", lambda: s)
    monkeypatch.setattr(launcher, "main", lambda argv=None: calls.append(argv) or 0)
    monkeypatch.setattr(worker, "run_worker", lambda: (_ for _ in ()).throw(AssertionError("served dp")))
    assert worker.main() == 0 and calls == [[]]
    s.gang_size = 1  # gangs of one = the plain worker
    assert not worker._wants_gangs(s)
    s.parallel_mode, s.gang_size = "dp", 8
    assert not worker._wants_gangs(s)
