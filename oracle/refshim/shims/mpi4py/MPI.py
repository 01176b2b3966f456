"""Single-process stand-in for mpi4py.MPI: just enough for ndsl.logging and imports."""


class Op:
    def __init__(self, name):
        self.name = name


SUM, MIN, MAX = Op("sum"), Op("min"), Op("max")


class Intracomm:
    def Get_rank(self):
        return 0

    def Get_size(self):
        return 1

    def bcast(self, v, root=0):
        return v

    def barrier(self):
        pass

    def allreduce(self, v, op=None):
        return v

    def Split(self, color=0, key=0):
        return self


Comm = Intracomm
COMM_WORLD = Intracomm()
