"""Known-answer linear least squares problems transcribed (data only) from the
reference's own test fixtures, /root/reference/internal/ceres/
linear_least_squares_problems.cc:
  problem 2  :301-423  (6x5, scalar blocks, 2 e-blocks, x to 4 d.p. at :311-316)
  problem 3  :425-527  (5x2, no f-blocks)
  problem 4  :528-626  (3x7, mixed f-block sizes, D = 100*(i+1))
  problem 5  :628-800  (6x5, block-diagonal F'F; S, r, S\\r, A\\b in the comment :628-676)
  problem 6  :802-929  (5x7, D = 100*(i+1))
Each entry: rows = [(row_size, row_pos, [(col_block_id, value_position), ...])],
cols = [(size, position)], values, b, D, num_eliminate_blocks, optional x.
"""
import numpy as np


def _scalar_cols(n):
    return [(1, c) for c in range(n)]


PROBLEMS = {
    2: dict(
        cols=_scalar_cols(5),
        rows=[(1, 0, [(0, 0), (2, 1)]), (1, 1, [(0, 2), (3, 3)]),
              (1, 2, [(1, 4), (4, 5)]), (1, 3, [(1, 6), (2, 7)]),
              (1, 4, [(1, 8), (2, 9)]), (1, 5, [(2, 10), (3, 11), (4, 12)])],
        values=[1, 2, 3, 4, 5, 6, 7, 8, 9, 1, 1, 1, 1],
        b=[0, 1, 2, 3, 4, 5], D=[1, 1, 1, 1, 1], num_eliminate_blocks=2,
        x=[-2.3061, 0.3172, 0.2102, 2.1367, 0.1388]),
    3: dict(
        cols=_scalar_cols(2),
        rows=[(1, 0, [(0, 0)]), (1, 1, [(0, 1)]), (1, 2, [(1, 2)]),
              (1, 3, [(1, 3)]), (1, 4, [(1, 4)])],
        values=[1, 3, 5, 7, 9],
        b=[0, 1, 2, 3, 4], D=[1, 1], num_eliminate_blocks=2),
    4: dict(
        cols=[(2, 0), (3, 2), (2, 5)],
        rows=[(2, 0, [(0, 0), (2, 4)]), (1, 2, [(1, 8), (2, 11)])],
        values=[1, 2, 1, 4, 1, 1, 5, 6, 9, 0, 0, 3, 1],
        b=[0, 1, 2], D=[100.0 * (i + 1) for i in range(7)],
        num_eliminate_blocks=1),
    5: dict(
        cols=_scalar_cols(5),
        rows=[(1, 0, [(0, 0), (4, 1)]), (1, 1, [(0, 2), (4, 3)]),
              (1, 2, [(1, 4), (3, 5)]), (1, 3, [(1, 6), (3, 7)]),
              (1, 4, [(1, 8), (2, 9)]), (1, 5, [(1, 10), (2, 11)])],
        values=[-1, 2, 3, 4, -1, 1, -3, 1, -1, 3, -2, 1],
        b=[0, 1, 2, 3, 4, 5], D=[1, 1, 1, 1, 1], num_eliminate_blocks=2,
        # the comment block :628-676 (for D = 0; note it states A(0,0)=1 while the
        # code stores -1; S, r and S\r below are unaffected by that sign)
        S=[[8.3333, -1.3333, 0.0], [-1.3333, 0.9333, 0.0], [0.0, 0.0, 10.0]],
        r=[8.6667, -1.6667, 1.0],
        S_solve_r=[0.9778, -0.3889, 0.1000]),
    6: dict(
        cols=[(2, 0), (3, 2), (2, 5)],
        rows=[(2, 0, [(0, 0), (2, 4)]), (2, 2, [(0, 8), (2, 12)]),
              (1, 4, [(1, 16), (2, 19)])],
        values=[1, 2, 1, 4, 1, 1, 5, 6, 3, 4, 5, 6, 7, 8, 9, 0, 9, 0, 0, 3, 1],
        b=[0, 1, 2, 3, 4], D=[100.0 * (i + 1) for i in range(7)],
        num_eliminate_blocks=1),
}


def flat(pid):
    """Return (structure kwargs for oracle.binding.Structure, values, b, D, extras)."""
    p = PROBLEMS[pid]
    row_size = [r[0] for r in p["rows"]]
    row_pos = [r[1] for r in p["rows"]]
    cell_ptr, cid, cpos = [0], [], []
    for r in p["rows"]:
        for (b, pos) in r[2]:
            cid.append(b)
            cpos.append(pos)
        cell_ptr.append(len(cid))
    st = dict(row_size=row_size, row_pos=row_pos, cell_ptr=cell_ptr,
              cell_block_id=cid, cell_position=cpos,
              col_size=[c[0] for c in p["cols"]], col_pos=[c[1] for c in p["cols"]],
              num_eliminate_blocks=p["num_eliminate_blocks"])
    return (st, np.array(p["values"], dtype=np.float64),
            np.array(p["b"], dtype=np.float64), np.array(p["D"], dtype=np.float64), p)
