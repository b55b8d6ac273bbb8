"""CPU: the Hydra-subset loader resolves the reference's unmodified task YAMLs to the dict the tasks read
(``isaacgymenvs/__init__.py:9-57``; SURVEY.md appendix C scalar table)."""
import handarm_b200.config as hcfg


def test_allegro_kuka_reorientation_scalars():
    cfg = hcfg.load_task_config("AllegroKuka", num_envs=32768, sim_device="cuda:0")
    env = cfg["env"]
    assert cfg["name"] == "AllegroKuka" and env["subtask"] == "reorientation"
    assert env["numEnvs"] == 32768                                  # ${resolve_default:8192,${...num_envs}}
    assert env["episodeLength"] == 600 and env["clampAbsObservations"] == 10.0
    assert env["successTolerance"] == 0.075 and env["targetSuccessTolerance"] == 0.01
    assert env["maxConsecutiveSuccesses"] == 50 and env["successSteps"] == 1
    assert env["liftingBonus"] == 300 and env["reachGoalBonus"] == 1000 and env["keypointScale"] == 1.5
    assert env["forceScale"] == 2.0 and env["forceProbRange"] == [0.001, 0.1]
    assert abs(cfg["sim"]["dt"] - 0.01667) < 1e-9 and cfg["sim"]["use_gpu_pipeline"] is True


def test_default_num_envs_and_subtask_group():
    cfg = hcfg.load_task_config("AllegroKuka")
    assert cfg["env"]["numEnvs"] == 8192                            # resolve_default falls back to the YAML default
    cfg = hcfg.load_task_config("AllegroKuka", subtask="regrasping")
    assert cfg["env"]["subtask"] == "regrasping" and cfg["env"]["successSteps"] == 30 and cfg["env"]["episodeLength"] == 300


def test_layered_two_arms_config():
    cfg = hcfg.load_task_config("AllegroKukaTwoArmsLSTM", num_envs=8192)
    assert cfg["name"] == "AllegroKukaTwoArms"                   # the LSTM file keeps the task class name
    assert cfg["env"]["armXOfs"] == 1.1 and cfg["env"]["envSpacing"] == 1.75
    assert cfg["env"]["episodeLength"] in (300, 600)


def test_overrides_and_other_hot_path_tasks_load():
    cfg = hcfg.load_task_config("AllegroKuka", num_envs=64, overrides={"env.successSteps": 30, "env.forceScale": 0.0})
    assert cfg["env"]["successSteps"] == 30 and cfg["env"]["forceScale"] == 0.0
    for name, n in (("ShadowHand", 16384), ("AllegroHand", 16384), ("AllegroHandDextremeADR", 16384)):
        c = hcfg.load_task_config(name)
        assert c["env"]["numEnvs"] == n and "sim" in c and "task" in c
    adr = hcfg.load_task_config("AllegroHandDextremeADR")
    assert adr["env"]["clipObservations"] == 50.0 and adr["task"]["adr"]["use_adr"] in (True, False)


def test_task_registry_dispatch():
    from handarm_b200.tasks import isaacgym_task_map
    assert "AllegroKuka" in isaacgym_task_map and "AllegroKukaLSTM" in isaacgym_task_map


def test_table_matches_reference_tree():
    """The shipped value table resolves to the same dicts as the reference's own YAML tree (checked where that tree exists)."""
    import os
    import pytest
    ref = "/root/reference/isaacgymenvs/cfg"
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present")
    cases = [(t, None) for t in hcfg.available_tasks()] + [("AllegroKuka", s) for s in ("regrasping", "throw")] + \
            [("AllegroKukaTwoArmsLSTM", "regrasping")]
    from_table = {c: hcfg.load_task_config(c[0], num_envs=77, subtask=c[1]) for c in cases}
    hcfg.set_cfg_dir(ref)
    try:
        assert set(hcfg.available_tasks()) >= {c[0] for c in cases}
        for c in cases:
            assert hcfg.load_task_config(c[0], num_envs=77, subtask=c[1]) == from_table[c], c
    finally:
        hcfg.set_cfg_dir(None)


def test_unknown_node_is_loud():
    import pytest
    with pytest.raises(KeyError):
        hcfg.load_task_config("NoSuchTask")
    with pytest.raises(FileNotFoundError):
        hcfg.set_cfg_dir("/nonexistent")
