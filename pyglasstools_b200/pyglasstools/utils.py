"""System data: simulation box and GSD trajectories (reference: pyglasstools/utils.py:8-141)."""
import os

import numpy as np

import pyglasstools
from pyglasstools import _pyglasstools, rank

try:                                    # the real package when it is installed ...
    import gsd.hoomd as _gsd_hoomd
except ImportError:                     # ... else our reader of the same file format
    from pyglasstools import gsdio as _gsd_hoomd


class simbox(object):
    """Orthorhombic periodic box; attribute access falls through to the C++ ``SimBox``
    (``dim, boxsize, origin, lowerbound, upperbound, periodic, vol``)."""

    def __init__(self, Lx=1.0, Ly=1.0, Lz=1.0, origin=np.zeros(3).astype('float64'), ndim=3, L=None, volume=None):
        if ndim == 2:
            Lz = 1.0
        if L is not None:
            self.cppsimbox = _pyglasstools.SimBox(float(L), origin, ndim)
        else:
            self.cppsimbox = _pyglasstools.SimBox(np.array([Lx, Ly, Lz], dtype=np.float64), origin, ndim)

    def __getattr__(self, attr):
        if attr == "cppsimbox":
            raise AttributeError(attr)
        return getattr(self.cppsimbox, attr)


def read_gsd(filename, checkpointfile=None, frame_num=0, ndim=2):
    traj = _gsd_hoomd.open(name=filename, mode='rb')
    return data_gsd(traj, checkpointfile, frame_num, ndim)


class data_gsd(object):
    """Particle data of one frame of a GSD trajectory, mirrored into a device-backed ``ParticleSystem``."""

    def __init__(self, traj_gsd, checkpointfile, frame_num, ndim=2):
        self.frame_num = frame_num
        self.checkpointfile = checkpointfile
        self.traj = traj_gsd
        self.ndim = ndim
        snap = self.traj[frame_num]
        box = snap.configuration.box
        self.pysimbox = simbox(Lx=box[0], Ly=box[1], Lz=box[2], ndim=ndim)
        self.cppparticledata = self._make_particles(snap)
        pyglasstools.set_sysdata(self)

    def _make_particles(self, snap):
        # NB the box of the FIRST loaded frame is used for every later frame, as in the reference
        # (utils.py:88-94 builds self.simbox but passes self.pysimbox.cppsimbox).
        p = snap.particles
        return _pyglasstools.ParticleSystem(self.pysimbox.cppsimbox, len(p.diameter), p.diameter, p.mass, p.position,
                                            p.velocity)

    def setup_checkpoint(self, frame_list):
        """Creates the checkpoint file, or resumes from the frame number it holds (utils.py:71-83)."""
        if self.checkpointfile is None:
            return frame_list
        if not os.path.exists(self.checkpointfile):
            if rank == 0:
                with open(self.checkpointfile, "w") as f:
                    f.write("Frame {}".format(frame_list[0]))
            return frame_list
        with open(self.checkpointfile, "r") as f:
            first = int(f.readline().split(' ')[1])
        return [i for i in range(first, frame_list[-1] + 1)]

    def update(self, frame_num):
        if frame_num != self.frame_num:
            snap = self.traj[frame_num]
            box = snap.configuration.box
            self.simbox = simbox(Lx=box[0], Ly=box[1], Lz=box[2], ndim=self.ndim)
            # the reference builds a new ParticleSystem per frame (utils.py:88-94); refilling the one at hand gives the same
            # particle data without 88 bytes per particle of fresh, page-faulting host memory per frame
            p = snap.particles
            self.cppparticledata.update(len(p.diameter), p.diameter, p.mass, p.position, p.velocity)
            self.frame_num = frame_num

    # thin forwards to the C++ object
    def move_particles(self, displacement=None):
        if displacement is None:
            self.cppparticledata.moveParticles()
        else:
            self.cppparticledata.moveParticles(displacement)

    def set_displacement(self, displacement):
        self.cppparticledata.setDisplacement(displacement)

    def get_displacement(self):
        return self.cppparticledata.getDisplacement()

    def set_diameters(self, diameter):
        self.cppparticledata.setDiameter(diameter.astype('float64'))

    def get_diameters(self):
        return self.cppparticledata.getDiameter()

    def set_mass(self, mass):
        self.cppparticledata.setMass(mass.astype('float64'))

    def get_mass(self):
        return self.cppparticledata.getMass()

    def set_position(self, position):
        self.cppparticledata.setAtomPosition(position.astype('float64'))

    def get_position(self):
        return self.cppparticledata.getAtomPosition()

    def set_velocity(self, velocity):
        self.cppparticledata.setAtomVelocity(velocity.astype('float64'))

    def get_velocity(self):
        return self.cppparticledata.getAtomVelocity()

    def get_neighbors(self, point, radius):
        return self.cppparticledata.getNeighbors(point, radius)

    def get_neighborsdistance(self, tag, radius):
        return self.cppparticledata.getNeighborsDistance(tag, radius)
