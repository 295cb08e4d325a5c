"""Forcing files of the `event_series` scenario (the file layer in front of the per-step grids).

Reference:
  src/scenario/ExternalForcing.cpp:29-51      a file = one forcing variable + `time` (values, `units`, `calendar`); next_timestep()
                                              reads the record at the cursor and returns its time value, -1 past the end
  src/scenario/EventSeriesScenario.cpp:71-119 record shapes (time, sector, region) for agents "SECTOR:REGION" / (time, sea_route) for
                                              geo entities; names that match nothing are skipped with a warning
  src/scenario/EventSeriesScenario.cpp:45-69  NaN -> forcing 1.0 (none); negative passage -> -1 (none)
  src/scenario/ExternalScenario.cpp:43-66     file name / expression templates: [[index]] and [[<key of scenario.parameters>]]
  src/scenario/ExternalScenario.cpp:68-83     reference year out of "days since Y-1-1" / "seconds since Y-1-1 00:00:00"
  src/scenario/ExternalScenario.cpp:85-127    next file: same calendar, units may move on by exactly one year (then the file's times
                                              count from the step after the current one), seconds are turned into days
  src/scenario/ExternalScenario.cpp:129-153   scenario.forcing.{variable, file, remove, index_from, index_to, expression}
  src/scenario/ExternalScenario.cpp:162-191   per step: a record is applied once the model time has reached its time value; when the last
                                              file is exhausted every agent's forcing goes back to 1.0

The reference reads NetCDF-4 through libnetcdf; HDF5 is not on this box, so files are NetCDF-3 classic (scipy.io.netcdf_file) with
the names as fixed-width character arrays -- the same variables, dimensions and attributes otherwise (`ncks -3` / `nccopy -k classic`
converts).  A record goes to the device as one array per step through Engine.set_agent_forcing / set_entity_passage, the same two
calls engine.EventSeriesScenario makes for in-memory grids.
"""
from __future__ import annotations

import os
import subprocess
import warnings

import numpy as np
from scipy.io import netcdf_file


class ForcingError(RuntimeError):
    pass


def _names(var):
    """A name variable of a classic file: (n, width) characters -> n strings."""
    a = np.asarray(var[:])
    if a.ndim == 1 and a.dtype.kind in "SU" and a.dtype.itemsize == 1:
        a = a.reshape(1, -1)
    if a.ndim != 2:
        raise ForcingError("name variables must be character arrays (name, width) in a classic NetCDF file")
    return [b"".join(row.tolist()).split(b"\0")[0].decode().strip() for row in a.astype("S1")]


class ForcingFile:
    """One forcing file (ExternalForcing + EventSeriesScenario::EventForcing)."""

    def __init__(self, filename, variable_name):
        if not os.path.exists(filename):
            raise ForcingError(f"Forcing file {filename} not found")
        self.filename = filename
        self.f = netcdf_file(filename, "r", mmap=False)
        if variable_name not in self.f.variables:
            raise ForcingError(f"Variable '{variable_name}' not found in {filename}")
        if "time" not in self.f.variables:
            raise ForcingError(f"Variable 'time' not found in {filename}")
        self.variable = self.f.variables[variable_name]
        self.time_variable = self.f.variables["time"]
        self.time_index_count = int(self.time_variable.shape[0])
        self.time_index = 0
        v = self.f.variables
        self.sectors = self.regions = self.sea_routes = None
        if "region" in v and "sector" in v:                      # EventSeriesScenario.cpp:81-103
            self.regions, self.sectors = _names(v["region"]), _names(v["sector"])
        if "sea_route" in v:                                     # :104-118
            self.sea_routes = _names(v["sea_route"])
        self.record = None

    def _attr(self, name):
        a = getattr(self.time_variable, name, None)
        if a is None:
            raise ForcingError(f"Attribute '{name}' of 'time' not found in {self.filename}")
        return a.decode() if isinstance(a, bytes) else str(a)

    def calendar_str(self):
        return self._attr("calendar")

    def time_units_str(self):
        return self._attr("units")

    def next_timestep(self):
        """ExternalForcing::next_timestep: loads the record at the cursor, returns its time value (an int), -1 at the end."""
        if self.time_index >= self.time_index_count:
            return -1
        self.record = np.array(self.variable[self.time_index], dtype=np.float64)   # EventForcing::read_data
        day = int(self.time_variable[self.time_index])
        self.time_index += 1
        return day

    def close(self):
        self.f.close()


def get_ref_year(filename, time_str):
    """ExternalScenario::get_ref_year (ExternalScenario.cpp:68-83), character positions included: for "seconds since " the reference
    looks one character past the '-' that follows the year (:76), so it accepts no such string -- kept as it is (it is only asked when
    the units change between two files)."""
    if time_str[:11] == "days since ":
        if time_str[15:19] != "-1-1" and time_str[15:21] != "-01-01":
            raise ForcingError(f"Forcing file {filename} has invalid time units")
        return int(time_str[11:15])
    if time_str[:14] == "seconds since ":
        if time_str[19:32] != "-1-1 00:00:00" and time_str[19:34] != "-01-01 00:00:00":
            raise ForcingError(f"Forcing file {filename} has invalid time units")
        return int(time_str[14:18])
    raise ForcingError(f"Forcing file {filename} has invalid time units")


class EventSeriesFiles:
    """scenario.type = event_series on forcing files: the stepping rule of ExternalScenario::iterate with the record rule of
    EventSeriesScenario::read_forcings.  `apply(engine, step)` is what ModelRun.step calls before the step's kernels."""

    def __init__(self, arrays, sector_names, region_names, scenario_node, entity_names=None, start_time=0.0, delta_t=1.0):
        self.node = scenario_node
        fn = scenario_node.get("forcing")
        if not isinstance(fn, dict) or "variable" not in fn or "file" not in fn:
            raise ForcingError("scenario.forcing.variable and scenario.forcing.file are required")   # :132-133
        for k in fn:
            if k not in ("variable", "file", "remove", "index_from", "index_to", "expression"):
                raise ForcingError(f"unknown key scenario.forcing.{k}")
        self.variable_name, self.forcing_file = str(fn["variable"]), str(fn["file"])
        self.remove_afterwards = bool(fn.get("remove", False))
        self.file_index_from = int(fn.get("index_from", 0))
        self.file_index_to = int(fn.get("index_to", self.file_index_from))
        self.file_index = self.file_index_from
        self.expression = str(fn["expression"]) if "expression" in fn else ""
        self.start_time, self.delta_t = float(start_time), float(delta_t)
        self.n_agents = int(arrays["meta"][2])
        asec, areg = np.asarray(arrays["agent_sector"]), np.asarray(arrays["agent_region"])
        self.agent_of = {(sector_names[int(s)], region_names[int(r)]): a for a, (s, r) in enumerate(zip(asec, areg))}
        self.sector_names = set(sector_names)
        self.entity_of = {n: i for i, n in enumerate(entity_names or [])}
        self.n_entities = len(entity_names or [])
        self.lam = np.ones(self.n_agents)
        self.nu = np.full(self.n_entities, -1.0)
        self.forcing = None
        self.calendar_str_ = self.time_units_str_ = ""
        self.time_step_width, self.time_offset, self.next_time = 1, 0.0, -1.0
        self.done = False
        self.model_time = 0.0     # ModelRun::run calls scenario->start() before it sets the model's start time (ModelRun.cpp:163-164)
        if not self._next_forcing_file():
            raise ForcingError("Empty forcing")                   # :149-151

    # ---- ExternalScenario ----
    def fill_template(self, text):
        out, pos = [], 0
        while True:
            start = text.find("[[", pos)
            stop = text.find("]]", start) if start >= 0 else -1
            if stop < 0:
                break
            out.append(text[pos:start])
            key = text[start + 2:stop]
            if key != "index":
                params = self.node.get("parameters", {})
                if key not in params:
                    raise ForcingError(f"scenario.parameters.{key} not found")
                out.append(str(params[key]))
            else:
                out.append(str(self.file_index))
            pos = stop + 2
        out.append(text[pos:])
        return "".join(out)

    def _open(self, filename):
        f = ForcingFile(filename, self.variable_name)
        f.agent_index = f.entity_index = None
        if f.sectors is not None:
            idx = np.full((len(f.sectors), len(f.regions)), -1, dtype=np.int64)
            for i, s in enumerate(f.sectors):
                if s not in self.sector_names:
                    warnings.warn(f"Sector '{s}' not found")
                for j, r in enumerate(f.regions):
                    a = self.agent_of.get((s, r), -1)
                    if s in self.sector_names and a < 0:
                        warnings.warn(f"Agent '{s}:{r}' not found")
                    idx[i, j] = a
            f.agent_index = idx
        if f.sea_routes is not None:
            idx = np.full(len(f.sea_routes), -1, dtype=np.int64)
            for i, n in enumerate(f.sea_routes):
                idx[i] = self.entity_of.get(n, -1)
                if idx[i] < 0:
                    warnings.warn(f"Sea object '{n}' not found")
            f.entity_index = idx
        return f

    def _next_forcing_file(self):
        if self.file_index > self.file_index_to:
            self._drop()
            return False
        if self.remove_afterwards and os.path.exists(self.forcing_file):
            os.remove(self.forcing_file)                          # the path as written in the settings (:90-92)
        if self.expression:
            cmd = self.fill_template(self.expression)
            if subprocess.call(cmd, shell=True) != 0:
                raise ForcingError(f"Invoking '{cmd}' raised an error")
        filename = self.fill_template(self.forcing_file)
        self._drop()
        self.forcing = self._open(filename)
        cal = self.forcing.calendar_str()
        if self.calendar_str_ and cal != self.calendar_str_:
            raise ForcingError("Forcing files differ in calendar")
        self.calendar_str_ = cal
        units = self.forcing.time_units_str()
        self.time_step_width = 24 * 60 * 60 if units[:14] == "seconds since " else 1
        if self.time_units_str_ and units != self.time_units_str_:
            ref_year = get_ref_year(filename, self.time_units_str_)
            if get_ref_year(filename, units) != ref_year + 1:
                raise ForcingError("Forcing files differ by more than a year")
            self.time_offset = self.model_time + self.delta_t
        self.time_units_str_ = units
        self.next_time = self.forcing.next_timestep() / self.time_step_width
        if self.next_time < 0:
            raise ForcingError(f"Empty forcing in {filename}")
        self.next_time += self.time_offset
        self.file_index += 1
        return True

    def _drop(self):
        if self.forcing is not None:
            self.forcing.close()
            self.forcing = None

    # ---- EventSeriesScenario::read_forcings ----
    def _read_forcings(self, engine):
        f = self.forcing
        rec = f.record
        if f.agent_index is not None:
            if rec.shape != f.agent_index.shape:
                raise ForcingError(f"{f.filename}: records must have the dimensions (sector, region)")
            ok = f.agent_index >= 0
            v = rec[ok]
            self.lam[f.agent_index[ok]] = np.where(np.isnan(v), 1.0, v)
            engine.set_agent_forcing(self.lam)
        elif f.entity_index is not None:
            if rec.shape != f.entity_index.shape:
                raise ForcingError(f"{f.filename}: records must have the dimension (sea_route)")
            ok = f.entity_index >= 0
            v = rec[ok]
            self.nu[f.entity_index[ok]] = np.where(v < 0.0, -1.0, v)
            engine.set_entity_passage(np.ascontiguousarray(self.nu))

    def apply(self, engine, step):
        """ExternalScenario::iterate at timestep `step` (model time = start + step * delta_t)."""
        if self.done:
            return
        self.model_time = self.start_time + step * self.delta_t
        if self.next_time < 0:
            if not self._next_forcing_file():
                self.done = True
                self.lam[:] = 1.0                                  # :175-181
                engine.set_agent_forcing(self.lam)
                return
        if self.model_time >= self.next_time:
            self._read_forcings(engine)
            self.next_time = self.forcing.next_timestep() / self.time_step_width
            if self.next_time >= 0:
                self.next_time += self.time_offset

    def close(self):
        """ExternalScenario::end"""
        self._drop()
        if self.remove_afterwards and os.path.exists(self.forcing_file):
            os.remove(self.forcing_file)


def write_forcing_file(path, variable, times, records, units="days since 2000-1-1", calendar="standard", sectors=None, regions=None,
                       sea_routes=None):
    """A classic NetCDF forcing file with the layout ForcingFile reads (fixtures, converters)."""
    records = np.asarray(records, dtype=np.float64)
    with netcdf_file(path, "w") as f:
        f.createDimension("time", len(times))
        t = f.createVariable("time", "i4", ("time",))
        t[:] = np.asarray(times, dtype=np.int32)
        t.units, t.calendar = units, calendar

        def names(dim, vals):
            w = max(len(s) for s in vals) + 1
            f.createDimension(dim, len(vals))
            f.createDimension(dim + "_len", w)
            v = f.createVariable(dim, "c", (dim, dim + "_len"))
            for i, s in enumerate(vals):
                v[i] = np.frombuffer(s.encode().ljust(w, b"\0"), dtype="S1")

        if sectors is not None:
            names("sector", sectors)
            names("region", regions)
            v = f.createVariable(variable, "f8", ("time", "sector", "region"))
        else:
            names("sea_route", sea_routes)
            v = f.createVariable(variable, "f8", ("time", "sea_route"))
        v[:] = records
