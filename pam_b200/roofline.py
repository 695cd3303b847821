"""Algorithmic work of the AWFL time step, frozen here so bench.py and DESIGN.md quote the same numbers.

Flop counts follow the REFERENCE's algorithm as written (one scalar operation = 1 flop, an FMA would be 2, a divide
counts as 1): weno::compute_weno_coefs (dynamics/awfl/WenoLimiter.h:98-181) + the edge evaluation in
Dycore::reconstruct (dynamics/awfl/Dycore.h:599-603), and 4 acoustic + 4 advective + NT tracer reconstructions per
interface and direction (Dycore.h:339-386). Bytes follow SURVEY.md 8(d): with ideal fusion one SSPRK3 sub-cycle reads
and writes each of the N = 5+NT fields 8 times in total.
"""

_TV_TERMS = {2: 1, 3: 2, 4: 4, 5: 6, 7: 12, 9: 20}   # terms of coefs_to_tv for n coefficients (tools/gen_tables.py)

# Instrumented count of one Dycore::reconstruct: the reference's algorithm, operation for operation, run on a scalar type
# that counts its additions/subtractions, multiplications and divisions (tools/flopcount/count_flops.cpp over
# pam_b200/cxx/Dycore.h reconstruct_generic; `python tools/count_flops.py`, output in profiles/r02_flop_count.json).
# FROZEN: tests/test_host_logic.py re-counts and compares. order -> (add, mul, div)
RECON_OPS = {3: (66, 67, 16), 5: (132, 141, 22), 7: (236, 261, 28), 9: (382, 429, 34)}


def recon_flops(ord):
    """(flops, of which divides) of one Dycore::reconstruct at order `ord` (instrumented count; a divide counts as 1)."""
    a, m, d = RECON_OPS[ord]
    return a + m + d, d


def recon_flops_hand_formula(ord):
    """The hand count the instrumented one replaced (kept as a cross-check: within 2 % at every order)."""
    m = (ord + 1) // 2
    tv = lambda n: 3 * _TV_TERMS[n] - 1
    add = 0
    add += m * m * 2 * m            # low-order polynomials: m*m outputs, m (mul+add) each
    add += ord * 2 * ord            # high-order polynomial
    add += 2 * m * m                # bridge: a_hi -= idl*a_lo
    add += m * tv(m) + tv(ord)      # total variations
    add += m + 3                    # lo_avg sum, damped bridge TV
    add += (m + 1) * 2              # tv*tv + eps
    add += 2 * 2 * (m + 1)          # two convexify: sum and sum+eps
    add += (m + 1) * 13             # map_weights
    add += ord + 2 * m * m          # blended coefficients
    add += 2 * ord                  # edge value
    div = ord + 1 + (m + 1) + 2 * (m + 1) + (m + 1)
    return add + div, div


def flops_per_cell_stage(ord, nt, ndir):
    """Reference flops per cell per Runge-Kutta stage that reach a result: reconstructions + everything else (products,
    characteristics, divergence, FCT, update, one pow counted as 1). A 2-D run (ndir == 2) has 3 advective reconstructions
    per direction, not 4: the reference reconstructs v and forms its flux there too, then zeroes the tendency
    (Dycore.h:565) -- work that no output depends on is not counted as algorithmic (it was until r02D, when the 2-D marches
    stopped executing it; counted, the state marches would read 1.00 of the FP64 peak)."""
    adv = 3 if ndir == 2 else 4
    per_dir = (4 + adv + nt) * recon_flops(ord)[0] + 2 * ord + 12 + 2 * (adv + nt) + 3 * (1 + adv + nt)
    other = 5 + 4 + nt + (5 + nt) * 5 + nt * 30
    return ndir * per_dir + other


def flops_per_cell_update(ord, nt, ndir):
    return 3 * flops_per_cell_stage(ord, nt, ndir)


def state_sweep_flops_per_cell(ord, sim2d=False, fused_tracer=False):
    """Reference flops of the part of the flux kernel the state sweep replaces, per cell and direction: 4 acoustic + 4
    advective reconstructions (3 in a 2-D run, see flops_per_cell_stage), stencil products, characteristics, the fluxes,
    divergence -- plus, where the march carries the first tracer, that tracer's share of the tracer march."""
    adv = 3 if sim2d else 4
    return (4 + adv) * recon_flops(ord)[0] + 2 * ord + 12 + 2 * adv + 3 * (1 + adv) + (tracer_sweep_flops_per_cell(ord) if fused_tracer else 0)


def tracer_sweep_flops_per_cell(ord):
    """Reference flops of one tracer's flux per cell and direction: one reconstruction, mixing ratio, upwind flux."""
    return recon_flops(ord)[0] + 4


def bytes_per_cell_update(nt):
    """Compulsory HBM bytes per cell per SSPRK3 sub-cycle with ideal fusion: 8*N doubles."""
    return 64 * (5 + nt)


def state_sweep_bytes_per_cell(last, fused_tracer=False):
    """Compulsory bytes of one state sweep launch per cell: read 5 conserved fields; a non-final sweep writes 5 partial
    tendencies + 1 mass flux; the final sweep reads 5 partial tendencies + 5 base fields, writes 5 fields + 1 mass flux. A
    march that carries the first tracer also reads that tracer and writes its flux."""
    return (8 * (5 + 5 + 1) if not last else 8 * (5 + 5 + 5 + 5 + 1)) + (16 if fused_tracer else 0)
