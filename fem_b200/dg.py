"""Command line of the scalar solver; same flags as dg1d/python/scalar/dg.py:10-53.

    python -m fem_b200.dg -pde burger -ic sin2pi -degree 2 -limit yes -num_flux roe -compute_error yes
"""
import argparse

import numpy as np

from .dgsolver import DG


def make_parser():
    p = argparse.ArgumentParser()
    p.add_argument('-pde', choices=('linear', 'varadv', 'burger'), help='PDE', default='linear')
    p.add_argument('-ncell', type=int, help='No of cells', default=50)
    p.add_argument('-nrefine', type=int, help='No of refinements', default=0)
    p.add_argument('-degree', type=int, help='Polynomial degree', default=1)
    p.add_argument('-cfl', type=float, help='CFL number', default=0.9)
    p.add_argument('-Tf', type=float, help='Final time', default=1.0)
    p.add_argument('-plot_freq', type=int, help='Accepted for compatibility; plotting is not built', default=1)
    p.add_argument('-ic', choices=('sin2pi', 'sin4pi', 'gauss', 'hat', 'mult'), help='Initial condition',
                   default='sin2pi')
    p.add_argument('-limit', choices=('no', 'yes'), help='Apply limiter', default='no')
    p.add_argument('-bc', choices=('periodic', 'dirichlet'), default='periodic',
                   help='Extension: dirichlet = inflow left / outflow right (dg1d/scalar_legendre/README.md:91-137)')
    p.add_argument('-tvbM', type=float, help='TVB M parameter', default=0.0)
    p.add_argument('-compute_error', choices=('no', 'yes'), help='Compute error norm', default='no')
    p.add_argument('-num_flux', choices=('central', 'upwind', 'roe', 'godunov'), help='Numerical flux',
                   default='upwind')
    return p


def main(argv=None):
    args = make_parser().parse_args(argv)
    if args.nrefine == 0:
        dg = DG(args)
        if args.compute_error == 'yes':
            nc, dx, error = dg.run()
            print('ncell, dx, error = %5d %12.4e %12.4e' % (nc, dx, error))
        else:
            dg.run()
        return dg
    # grid refinement study, dg.py:39-53
    args.plot_freq = 0
    args.compute_error = 'yes'
    error = np.empty(args.nrefine)
    for n in range(args.nrefine):
        dg = DG(args)
        nc, dx, error[n] = dg.run()
        if n == 0:
            print('ncell, dx, error, rate = %5d %12.4e %12.4e' % (nc, dx, error[n]))
        else:
            rate = np.log(error[n - 1] / error[n]) / np.log(2)
            print('ncell, dx, error, rate = %5d %12.4e %12.4e %8.3f' % (nc, dx, error[n], rate))
        args.ncell *= 2
    return error


if __name__ == '__main__':
    main()
