"""Task registry (subset of ``isaacgymenvs/tasks/__init__.py:66-123`` that lies on the hot path)."""
from .allegro_kuka import (AllegroKukaBase, AllegroKukaRegrasping, AllegroKukaReorientation, AllegroKukaThrow,
                           AllegroKukaTwoArmsRegrasping, AllegroKukaTwoArmsReorientation)
from .dextreme import AllegroHandDextremeADR
from .hand import AllegroHand, ShadowHand


def resolve_allegro_kuka(cfg, *args, **kwargs):
    """``tasks/__init__.py:66-76``: dispatch on ``env.subtask``."""
    subtask = cfg["env"]["subtask"]
    table = {"reorientation": AllegroKukaReorientation, "regrasping": AllegroKukaRegrasping, "throw": AllegroKukaThrow}
    if subtask not in table:
        raise ValueError(f"Unknown subtask={subtask} in {list(table)}")
    return table[subtask](cfg, *args, **kwargs)


def resolve_allegro_kuka_two_arms(cfg, *args, **kwargs):
    """``tasks/__init__.py:80-90``."""
    subtask = cfg["env"]["subtask"]
    table = {"regrasping": AllegroKukaTwoArmsRegrasping, "reorientation": AllegroKukaTwoArmsReorientation}
    if subtask not in table:
        raise ValueError(f"Unknown subtask={subtask} in {list(table)}")
    return table[subtask](cfg, *args, **kwargs)


isaacgym_task_map = {"AllegroKukaTwoArms": resolve_allegro_kuka_two_arms, "AllegroKukaTwoArmsLSTM": resolve_allegro_kuka_two_arms,
                     "AllegroKuka": resolve_allegro_kuka, "AllegroKukaLSTM": resolve_allegro_kuka, "ShadowHand": ShadowHand,
                     "AllegroHand": AllegroHand, "AllegroHandDextremeADR": AllegroHandDextremeADR}
