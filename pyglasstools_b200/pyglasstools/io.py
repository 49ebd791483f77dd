"""Loggers (reference: pyglasstools/io.py:9-169): ``logfile`` writes one text line of global observables per
frame, ``fieldlogger`` dumps coarse-grained fields as ``<keyword>.xyz``.  Formats are byte-compatible with the
reference: ``%f`` globals, 6-significant-digit field values, ``str(float)`` coordinates."""
import os

import pyglasstools
from pyglasstools import _pyglasstools, comm, irvingkirkwood, loggers_list, rank, size, thermo


def _require(condition, message):
    if not condition:
        if rank == 0:
            print("[ERROR] " + message)
        comm.abort(1)


def _checkpoint_is_fresh():
    path = pyglasstools.get_sysdata().checkpointfile
    return not os.path.exists(path) or os.path.getsize(path) == 0


def _tensor_index(name, dim, rank_):
    """Flat index of the trailing digits of ``name``: 'ab' -> a + dim b (col-major);
    'abcd' -> d + dim (c + dim (b + dim a)) as the reference's logger does for the Born tensor."""
    digits = [int(c) for c in name[-rank_:]]
    if rank_ == 2:
        return digits[0] + dim * digits[1]
    i, j, k, l = digits
    return l + dim * (k + dim * (j + dim * i))


class logfile(object):

    def __init__(self, filename, names, solver):
        self.filename = filename
        self.__obs_names = names
        _require(solver is not None, "User must specify solver object!")
        _require(names is not None, "User must specify list of observables to save!")
        _require(filename is not None, "User must specify a filename for the logfile!")
        self.solver = solver
        loggers_list.append(self)
        self.file = _pyglasstools.MPILogFile(comm, self.filename)
        self.global_obs = thermo.initialize_global(self.__obs_names, pyglasstools.get_sysdata().pysimbox.dim)
        self.solver.add_observables(self.global_obs)
        if rank == 0 and _checkpoint_is_fresh():
            self.file.write_shared("Frame " + "".join("{} ".format(n) for n in self.__obs_names) + "\n")

    def save(self, frame_num):
        if rank != 0:
            return
        dim = pyglasstools.get_sysdata().pysimbox.dim
        write = self.file.write_shared
        write("{} ".format(frame_num))
        for name in self.__obs_names:
            if "potentialenergy" in name:
                self.global_obs['potentialenergy'].save(self.file, 0)
                write(" ")
            if "virialstress" in name:
                self.global_obs['virialstress'].save(self.file, _tensor_index(name, dim, 2))
                write(" ")
            elif "kineticstress" in name:
                self.global_obs['kineticstress'].save(self.file, _tensor_index(name, dim, 2))
                write(" ")
            elif "borntensor" in name:
                self.global_obs['borntensor'].save(self.file, _tensor_index(name, dim, 4))
                write(" ")
        write("\n")
        self.file.flush()


class fieldlogger(object):

    def __init__(self, keyword, names, solver, binary=False):
        """``binary=True`` (additive extension, off by default): besides the reference's text file, every frame's
        requested columns are saved at full precision as ``<keyword>.frame<N>[.rank<r>].npy`` (rows = this rank's grid
        points, columns = x, y[, z], then the requested components in the order of ``names``)."""
        self.__obs_names = names
        _require(solver is not None, "User must specify solver object!")
        _require(names is not None, "User must specify list of observables to save!")
        self.solver = solver
        self.keyword = keyword
        self.binary = bool(binary)
        loggers_list.append(self)
        self.field_obs = irvingkirkwood.initialize_field(self.__obs_names, pyglasstools.get_sysdata().pysimbox.dim,
                                                         self.solver.gridpoints)
        self.solver.add_observables(self.field_obs)
        self.file = _pyglasstools.MPILogFile(comm, "{}".format(keyword) + ".xyz")

    def _columns(self, dim):
        fields, comps = [], []
        for name in self.__obs_names:
            fields.append(self.field_obs[name[:-3]])
            comps.append(_tensor_index(name, dim, 2))
        return fields, comps

    def save_perrank(self, frame_num):
        """All lines of this rank's chunk, formatted in one C++ pass (same bytes as the reference's per-point loop)."""
        dim = pyglasstools.get_sysdata().pysimbox.dim
        fields, comps = self._columns(dim)
        _pyglasstools.write_field_frame(self.file, self.solver.gridpoints, dim, fields, comps)
        if self.binary:
            import numpy as np
            pts = np.asarray(self.solver.gridpoints, dtype=np.float64)[:, :(3 if dim == 3 else 2)]
            cols = [f.values()[:, c] for f, c in zip(fields, comps)]
            suffix = "" if size == 1 else ".rank{}".format(rank)
            np.save("{}.frame{}{}.npy".format(self.keyword, frame_num, suffix), np.column_stack([pts] + cols))

    def save_perrank_polar(self, frame_num):
        dim = pyglasstools.get_sysdata().pysimbox.dim
        fields, comps = self._columns(dim)
        for index in range(len(self.solver.gridpoints)):
            outline = "1 {} {} {} ".format(int(self.solver.gridid[index]), self.solver.globalr[index],
                                           self.solver.globaltheta[index])
            for field, comp in zip(fields, comps):
                outline += field.gettostring(comp, index)
            self.file.write_shared(outline + "\n")

    def save(self, frame_num):
        if rank == 0:
            self.file.write_shared("{:d} \n".format(self.solver.gridsize))
            self.file.write_shared("#Frame {:d}  \n".format(frame_num))
            self.file.flush()
        comm.barrier()
        if self.solver.coordmode == "cartesian":
            self.save_perrank(frame_num)
        elif self.solver.coordmode == "polar":
            self.save_perrank_polar(frame_num)
        self.file.flush()
