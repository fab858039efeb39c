// Compiles the reference's own Roe flux -- PDE::roe_flux<dim> in dg2d/models/euler/pde.h:209-285, the function the oracle's
// orc_euler_roe_flux restates for one space dimension -- from the header where it lies, unmodified, against the deal.II
// stand-in types in oracle/dealii_standin (deal.II is not installed here).  TEST INFRASTRUCTURE, output oracle/_ref only.
// One space dimension = dim 2 with normal (1,0) and zero y-momentum: every y term is then an exact zero (0*0, x*1, x+0),
// so components 0, 1 and 3 of the 2-D flux are the 1-D flux bit for bit, whatever the summation order of the stand-ins.
#include "pde.h"

namespace ProblemData {
extern const double gamma = 1.4;  // pde.h:34-37 expects the problem to define it; dg1d/python/euler/gas.py:3 uses 1.4
}

extern "C" void ref_dg2d_roe_flux(const double ul[3], const double ur[3], double f[3])
{
    dealii::Vector<double> a(nvar), b(nvar), fl(nvar);
    a[0] = ul[0]; a[1] = ul[1]; a[2] = 0.0; a[3] = ul[2];
    b[0] = ur[0]; b[1] = ur[1]; b[2] = 0.0; b[3] = ur[2];
    dealii::Tensor<1, 2> n;
    n[0] = 1.0; n[1] = 0.0;
    PDE::roe_flux<2>(a, b, n, fl);
    f[0] = fl[0]; f[1] = fl[1]; f[2] = fl[3];
}
