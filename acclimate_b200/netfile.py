"""Flat "ACCARR1" container: 8-byte magic, then records
``u32 name_len | name | u8 dtype (0 int64, 1 float64) | u64 count | count*8 bytes``.

It carries the network snapshot (what the reference's ModelInitializer produces, flattened:
reference src/input/ModelInitializer.cpp:94-213,689-700) and the per-step dumps of the oracle.
"""
from __future__ import annotations

import gzip
import struct
from collections import OrderedDict

import numpy as np

MAGIC = b"ACCARR1\0"

# order of "model_f64" (mirrors reference include/parameters.h:33-84 for the keys this path reads)
MODEL_PARAMS = [
    "transport_penalty_small",
    "transport_penalty_large",
    "min_storage",
    "cheapest_price_range_width",  # < 0 means "auto"
    "optimization_maxiter",
    "optimization_precision_adjustment",
    "quadratic_transport_penalty",
    "relative_transport_penalty",
    "maximal_decrease_reservation_price_limited_by_markup",
    "deviation_penalty",
    "cheapest_price_range_preserve_seller_price",
    "no_self_supply",
    "utilitarian",
    "delta_t",
    "naive_expectations",
    "always_extend_expected_demand_curve",
]
# order of each row of "sector_f64" (reference ModelInitializer.cpp:159-185)
SECTOR_PARAMS = [
    "upper_storage_limit",
    "initial_storage_fill_factor",
    "transport_type",  # 0 aviation, 1 immediate, 2 roadsea
    "supply_elasticity",
    "price_increase_production_extension",
    "estimated_price_increase_production_extension",
    "initial_markup",
    "target_storage_refill_time",
    "target_storage_withdraw_time",
]


def _open(path, mode):
    """Files named *.gz are gzip-compressed (the golden fixtures); the oracle only ever reads plain files."""
    if str(path).endswith(".gz"):
        return gzip.GzipFile(path, mode, compresslevel=6, mtime=0) if "w" in mode else gzip.open(path, mode)
    return open(path, mode)


def write(path, arrays):
    with _open(path, "wb") as fp:
        fp.write(MAGIC)
        for name, arr in arrays.items():
            arr = np.ascontiguousarray(arr)
            if arr.dtype.kind in "iub":
                arr = arr.astype("<i8")
                dt = 0
            else:
                arr = arr.astype("<f8")
                dt = 1
            nb = name.encode()
            fp.write(struct.pack("<I", len(nb)))
            fp.write(nb)
            fp.write(struct.pack("<BQ", dt, arr.size))
            fp.write(arr.tobytes())


def read(path, prefix=None):
    out = OrderedDict()
    with _open(path, "rb") as fp:
        if fp.read(8) != MAGIC:
            raise ValueError(f"{path}: not an ACCARR1 file")
        while True:
            h = fp.read(4)
            if len(h) < 4:
                break
            (nl,) = struct.unpack("<I", h)
            name = fp.read(nl).decode()
            dt, cnt = struct.unpack("<BQ", fp.read(9))
            if prefix is not None and not name.startswith(prefix):
                fp.seek(cnt * 8, 1)
                continue
            out[name] = np.frombuffer(fp.read(cnt * 8), dtype="<i8" if dt == 0 else "<f8")
    return out
