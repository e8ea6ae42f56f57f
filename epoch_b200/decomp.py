"""Cartesian domain decomposition of the reference (host logic, pure Python).

Restates `split_domain` / `setup_communicator` / `mpi_initialise`
(epoch3d/src/housekeeping/mpi_routines.F90:44-593): rank <-> coordinates of the
MPI_CART_CREATE communicator with dims (nprocz, nprocy, nprocx) (x fastest),
near-even cell split, and the neighbour table used by field_bc / particle_bcs.
"""


def rank_to_coords(rank, nprocdir):
    """MPI_CART_COORDS for dims=(nprocz,nprocy,nprocx), row major -> (x,y,z)."""
    npx, npy, npz = (list(nprocdir) + [1, 1, 1])[:3]
    return (rank % npx, (rank // npx) % npy, rank // (npx * npy))


def coords_to_rank(coords, nprocdir):
    npx, npy, npz = (list(nprocdir) + [1, 1, 1])[:3]
    x, y, z = (list(coords) + [0, 0, 0])[:3]
    return (z * npy + y) * npx + x


def cell_range(n_global, nproc_dir, coord):
    """1-based inclusive [cell_min, cell_max] (mpi_routines.F90:444-476): the
    first nxp processors get nx0 cells, the remaining ones nx0+1."""
    nx0 = n_global // nproc_dir
    if nx0 * nproc_dir != n_global:
        nxp = (nx0 + 1) * nproc_dir - n_global
    else:
        nxp = nproc_dir
    idim = coord + 1
    if idim <= nxp:
        return (idim - 1) * nx0 + 1, idim * nx0
    return nxp * nx0 + (idim - nxp - 1) * (nx0 + 1) + 1, nxp * nx0 + (idim - nxp) * (nx0 + 1)


def neighbour(coords, nprocdir, axis, direction, periodic):
    """MPI_CART_SHIFT along `axis` (0=x) by `direction` (+-1); None == MPI_PROC_NULL."""
    c = list((list(coords) + [0, 0, 0])[:3])
    n = (list(nprocdir) + [1, 1, 1])[:3]
    c[axis] += direction
    if c[axis] < 0 or c[axis] >= n[axis]:
        if not periodic:
            return None
        c[axis] %= n[axis]
    return coords_to_rank(c, n)


def slab_decomposition(nproc, ndims):
    """The bench's choice: split the slowest existing axis only (contiguous halo
    slabs)."""
    dirs = [1, 1, 1]
    dirs[ndims - 1] = nproc
    return tuple(dirs)
