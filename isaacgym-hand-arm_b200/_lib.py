"""ctypes binding of ``libhandarm_b200.so`` (C-ABI declared in ``include/handarm_b200.h``).

The ctypes structures are generated from the header itself, so the Python side can never drift
from the C side; struct sizes are cross-checked against ``ha_sizeof_*`` at load time.

There is no fallback: if the shared library is missing or was built for a different ABI the import
of the compute path raises ``HandArmLibraryError``.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List, Tuple

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_REPO_DIR = os.path.dirname(_PKG_DIR)
HEADER_PATH = os.path.join(_REPO_DIR, "include", "handarm_b200.h")
LIB_PATH = os.environ.get("HANDARM_B200_LIB") or os.path.join(_PKG_DIR, "libhandarm_b200.so")   # env: tuning hook for A/B builds


class HandArmLibraryError(RuntimeError):
    pass


class HandArmError(RuntimeError):
    """Raised when a C-ABI call returns a negative status."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"handarm_b200 error {code}: {msg}")
        self.code = code


_CTYPES = {
    "int32_t": ctypes.c_int32,
    "uint32_t": ctypes.c_uint32,
    "int64_t": ctypes.c_int64,
    "uint64_t": ctypes.c_uint64,
    "uint8_t": ctypes.c_uint8,
    "float": ctypes.c_float,
    "int": ctypes.c_int,
    "size_t": ctypes.c_size_t,
}


def _strip_comments(src: str) -> str:
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def _parse_header(path: str):
    src = _strip_comments(open(path).read())
    consts: Dict[str, int] = {}
    for m in re.finditer(r"#define\s+(HA_\w+)\s+\(?(-?\d+)\)?\s*$", src, flags=re.M):
        consts[m.group(1)] = int(m.group(2))
    # enums (anonymous): sequential values with optional "= n"
    for m in re.finditer(r"enum\s*\{(.*?)\}\s*;", src, flags=re.S):
        val = -1
        for item in m.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                name, expr = [t.strip() for t in item.split("=")]
                val = int(eval(expr, {}, dict(consts)))
            else:
                name, val = item, val + 1
            consts[name] = val
    structs: Dict[str, List[Tuple[str, object]]] = {}
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s*\{(.*?)\}\s*(\w+)\s*;", src, flags=re.S):
        name, body = m.group(3), m.group(2)
        fields: List[Tuple[str, object]] = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            is_const = decl.startswith("const ")
            if is_const:
                decl = decl[len("const "):]
            tname, rest = decl.split(" ", 1)
            ptr = False
            if tname.endswith("*"):
                tname, ptr = tname[:-1], True
            for var in rest.split(","):
                var = var.strip()
                vptr = ptr
                while var.startswith("*"):
                    var, vptr = var[1:].strip(), True
                dims = [int(eval(d, {}, dict(consts))) for d in re.findall(r"\[([^\]]+)\]", var)]
                var = var.split("[")[0].strip()
                if vptr:
                    ctype = ctypes.c_void_p
                else:
                    ctype = _CTYPES[tname] if tname in _CTYPES else structs_ctypes[tname]
                for d in reversed(dims):
                    ctype = ctype * d
                fields.append((var, ctype))
        cls = type(name, (ctypes.Structure,), {"_fields_": fields})
        structs_ctypes[name] = cls
        structs[name] = fields
    return consts, structs_ctypes


structs_ctypes: Dict[str, type] = {}
CONSTS, STRUCTS = _parse_header(HEADER_PATH)

HaKukaParams = STRUCTS["HaKukaParams"]
HaKukaBuffers = STRUCTS["HaKukaBuffers"]
HaKukaDynamic = STRUCTS["HaKukaDynamic"]
HaHandParams = STRUCTS["HaHandParams"]
HaHandBuffers = STRUCTS["HaHandBuffers"]
HaHandDynamic = STRUCTS["HaHandDynamic"]
HaDexParams = STRUCTS["HaDexParams"]
HaDexBuffers = STRUCTS["HaDexBuffers"]
HaDexAdrTable = STRUCTS["HaDexAdrTable"]
HaDexAdrBuffers = STRUCTS["HaDexAdrBuffers"]
HaNoiseParams = STRUCTS["HaNoiseParams"]
HaObserverBuffers = STRUCTS["HaObserverBuffers"]
HaAdrQueueBuffers = STRUCTS["HaAdrQueueBuffers"]
HaStatsExchange = STRUCTS["HaStatsExchange"]

_lib = None


def exported_functions() -> List[str]:
    """Names of every function the header declares (used by the CPU-side ABI test)."""
    src = _strip_comments(open(HEADER_PATH).read())
    return sorted(set(re.findall(r"\b(ha_\w+)\s*\(", src)) - {"ha_stream_t"})


def load():
    """Load the shared library (once). Fails loudly when it is absent or ABI-incompatible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HandArmLibraryError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C isaacgym-hand-arm_b200/csrc`). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    lib.ha_last_error.restype = ctypes.c_char_p
    for fn in ("ha_sizeof_kuka_params", "ha_sizeof_kuka_buffers", "ha_sizeof_kuka_dynamic", "ha_kuka_bytes_per_env_step",
               "ha_kuka_post_bytes_per_env_step", "ha_kuka_pre_bytes_per_env_step"):
        getattr(lib, fn).restype = ctypes.c_size_t
    lib.ha_kuka_stat_accum_init.restype = ctypes.c_int64
    lib.ha_kuka_stat_accum_init.argtypes = [ctypes.c_int]
    if lib.ha_abi_version() != CONSTS["HA_ABI_VERSION"]:
        raise HandArmLibraryError(f"ABI mismatch: library {lib.ha_abi_version()} vs header {CONSTS['HA_ABI_VERSION']}")
    for cls, fn in ((HaKukaParams, lib.ha_sizeof_kuka_params), (HaKukaBuffers, lib.ha_sizeof_kuka_buffers),
                    (HaKukaDynamic, lib.ha_sizeof_kuka_dynamic)):
        if ctypes.sizeof(cls) != fn():
            raise HandArmLibraryError(f"struct {cls.__name__}: ctypes size {ctypes.sizeof(cls)} != C size {fn()}")
    lib.ha_kuka_pre_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_kuka_post_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_kuka_observe.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_kuka_stats_finalize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_kuka_reset_idx.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_kuka_obs_dim.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.ha_kuka_fill_derived.argtypes = [ctypes.c_void_p]
    lib.ha_kuka_fill_derived.restype = None
    for fn in ("ha_kuka_bytes_per_env_step", "ha_kuka_post_bytes_per_env_step", "ha_kuka_pre_bytes_per_env_step"):
        getattr(lib, fn).argtypes = [ctypes.c_void_p]
    # in-hand family
    for fn in ("ha_sizeof_hand_params", "ha_sizeof_hand_buffers", "ha_hand_bytes_per_env_step", "ha_hand_post_bytes_per_env_step",
               "ha_hand_pre_bytes_per_env_step"):
        getattr(lib, fn).restype = ctypes.c_size_t
    for cls, fn in ((HaHandParams, lib.ha_sizeof_hand_params), (HaHandBuffers, lib.ha_sizeof_hand_buffers)):
        if ctypes.sizeof(cls) != fn():
            raise HandArmLibraryError(f"struct {cls.__name__}: ctypes size {ctypes.sizeof(cls)} != C size {fn()}")
    lib.ha_hand_pre_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_hand_reset_idx.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_hand_post_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_hand_observe.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_hand_stats_finalize.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_hand_obs_dim.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.ha_hand_fill_derived.argtypes = [ctypes.c_void_p]
    lib.ha_hand_fill_derived.restype = None
    for fn in ("ha_hand_bytes_per_env_step", "ha_hand_post_bytes_per_env_step", "ha_hand_pre_bytes_per_env_step"):
        getattr(lib, fn).argtypes = [ctypes.c_void_p]
    # DeXtreme family
    for fn in ("ha_sizeof_dex_params", "ha_sizeof_dex_buffers", "ha_sizeof_dex_adr_table", "ha_sizeof_dex_adr_buffers", "ha_dex_bytes_per_env_step", "ha_dex_post_bytes_per_env_step",
               "ha_dex_pre_bytes_per_env_step"):
        getattr(lib, fn).restype = ctypes.c_size_t
    for cls, fn in ((HaDexParams, lib.ha_sizeof_dex_params), (HaDexBuffers, lib.ha_sizeof_dex_buffers),
                    (HaDexAdrTable, lib.ha_sizeof_dex_adr_table), (HaDexAdrBuffers, lib.ha_sizeof_dex_adr_buffers)):
        if ctypes.sizeof(cls) != fn():
            raise HandArmLibraryError(f"struct {cls.__name__}: ctypes size {ctypes.sizeof(cls)} != C size {fn()}")
    lib.ha_dex_pre_step.argtypes = [ctypes.c_void_p] * 6
    lib.ha_dex_post_step.argtypes = [ctypes.c_void_p] * 4
    lib.ha_dex_observe.argtypes = [ctypes.c_void_p] * 4
    lib.ha_dex_adr_resample.argtypes = [ctypes.c_void_p] * 6
    lib.ha_dex_fill_derived.argtypes = [ctypes.c_void_p]
    lib.ha_dex_fill_derived.restype = None
    for fn in ("ha_dex_bytes_per_env_step", "ha_dex_post_bytes_per_env_step", "ha_dex_pre_bytes_per_env_step"):
        getattr(lib, fn).argtypes = [ctypes.c_void_p]
    # DR noise
    lib.ha_sizeof_noise_params.restype = ctypes.c_size_t
    if ctypes.sizeof(HaNoiseParams) != lib.ha_sizeof_noise_params():
        raise HandArmLibraryError(f"struct HaNoiseParams: ctypes size {ctypes.sizeof(HaNoiseParams)} != C size {lib.ha_sizeof_noise_params()}")
    lib.ha_dr_noise_apply.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                      ctypes.c_void_p]
    # rl_games observer
    lib.ha_sizeof_observer_buffers.restype = ctypes.c_size_t
    if ctypes.sizeof(HaObserverBuffers) != lib.ha_sizeof_observer_buffers():
        raise HandArmLibraryError(f"struct HaObserverBuffers: ctypes size {ctypes.sizeof(HaObserverBuffers)} != C size {lib.ha_sizeof_observer_buffers()}")
    lib.ha_observer_step.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_observer_summary.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
    # random network adversary
    lib.ha_rna_input_layer.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 6
    lib.ha_rna_input_planes.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_rna_split_bf16.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.ha_rna_layer_tc.argtypes = [ctypes.c_int32] * 3 + [ctypes.c_void_p] * 6 + [ctypes.c_int32] + [ctypes.c_void_p] * 3
    lib.ha_rna_select_tc.argtypes = [ctypes.c_int32] * 2 + [ctypes.c_void_p] * 9
    lib.ha_rna_bias_act.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_size_t, ctypes.c_int32, ctypes.c_void_p]
    lib.ha_rna_select.argtypes = [ctypes.c_int32] + [ctypes.c_void_p] * 6
    # ADR objective queues
    lib.ha_sizeof_adr_queue_buffers.restype = ctypes.c_size_t
    if ctypes.sizeof(HaAdrQueueBuffers) != lib.ha_sizeof_adr_queue_buffers():
        raise HandArmLibraryError("struct HaAdrQueueBuffers: ctypes size != C size")
    lib.ha_adr_queue_step.argtypes = [ctypes.c_int32] * 4 + [ctypes.c_void_p, ctypes.c_void_p]
    lib.ha_sizeof_stats_exchange.restype = ctypes.c_size_t
    if ctypes.sizeof(HaStatsExchange) != lib.ha_sizeof_stats_exchange():
        raise HandArmLibraryError("struct HaStatsExchange: ctypes size != C size")
    lib.ha_stats_exchange_bytes.restype = ctypes.c_size_t
    lib.ha_stats_exchange_bytes.argtypes = [ctypes.c_int32]
    lib.ha_stats_exchange.argtypes = [ctypes.c_void_p] * 4
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise HandArmError(rc, load().ha_last_error().decode())
