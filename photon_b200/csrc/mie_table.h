// mie_table.h -- Bohren & Huffman Mie scattering amplitudes (host, double precision).
// Replaces python_codes/bhmie.py (called per particle diameter by run_simulation_02.py:641-696 to build the
// scattering_irradiance[angle][diameter] table that start_ray_tracing receives in scattering_data_t).
// Algorithm: logarithmic derivative D_n(mx) by downward recurrence from nmx = max(xstop, |mx|) + 15,
// Riccati-Bessel psi_n, chi_n by upward recurrence, a_n, b_n, angular functions pi_n, tau_n by upward
// recurrence; S1, S2 for nang angles in [0, 90] deg and their mirror images in (90, 180].
#pragma once
#include <cmath>
#include <complex>
#include <vector>

namespace ph {

// s1, s2: 2*nang-1 complex amplitudes (interleaved re, im), q: {qext, qsca, qback, gsca}.  Returns 0, or 1
// for the argument errors bhmie.py prints and returns on (nang > 1000, nmx > 150000).
inline int bhmie_host(double x, double m_re, double m_im, int nang, double *s1, double *s2, double *q) {
    typedef std::complex<double> cd;
    if (nang > 1000) return 1;
    if (nang < 2) nang = 2;
    const cd m(m_re, m_im);
    const cd y = x * m;
    const double ymod = std::abs(y);
    const double xstop = x + 4.0 * std::pow(x, 0.3333) + 2.0;
    const double nmx = std::trunc(std::max(xstop, ymod) + 15.0);
    const int nstop = (int)xstop;
    if (nmx > 150000.0) return 1;
    const double dang = 0.5 * (4.0 * std::atan(1.0)) / (nang - 1);
    std::vector<double> amu(nang), pi0(nang, 0.0), pi1(nang, 1.0), pi(nang), tau(nang);
    for (int j = 0; j < nang; j++) amu[j] = std::cos(j * dang);
    const int nn = (int)nmx - 1;
    std::vector<cd> d(nn + 1, cd(0.0, 0.0));
    for (int n = 0; n < nn; n++) {
        const double en = nmx - n;
        d[nn - n - 1] = (en / y) - (1.0 / (d[nn - n] + en / y));
    }
    std::vector<cd> s1a(nang, cd(0, 0)), s1b(nang, cd(0, 0)), s2a(nang, cd(0, 0)), s2b(nang, cd(0, 0));
    double psi0 = std::cos(x), psi1 = std::sin(x), chi0 = -std::sin(x), chi1 = std::cos(x);
    cd xi1(psi1, -chi1);
    double qsca = 0.0, gsca = 0.0, p = -1.0;
    cd an(0, 0), bn(0, 0), an1(0, 0), bn1(0, 0);
    for (int n = 0; n < nstop; n++) {
        const double en = n + 1.0;
        const double fn = (2.0 * en + 1.0) / (en * (en + 1.0));
        const double psi = (2.0 * en - 1.0) * psi1 / x - psi0;
        const double chi = (2.0 * en - 1.0) * chi1 / x - chi0;
        const cd xi(psi, -chi);
        if (n > 0) { an1 = an; bn1 = bn; }
        an = (d[n] / m + en / x) * psi - psi1;
        an = an / ((d[n] / m + en / x) * xi - xi1);
        bn = (m * d[n] + en / x) * psi - psi1;
        bn = bn / ((m * d[n] + en / x) * xi - xi1);
        qsca += (2.0 * en + 1.0) * (std::norm(an) + std::norm(bn));
        gsca += ((2.0 * en + 1.0) / (en * (en + 1.0))) * (an.real() * bn.real() + an.imag() * bn.imag());
        if (n > 0)
            gsca += ((en - 1.0) * (en + 1.0) / en) * (an1.real() * an.real() + an1.imag() * an.imag() +
                                                      bn1.real() * bn.real() + bn1.imag() * bn.imag());
        p = -p;
        for (int j = 0; j < nang; j++) {
            pi[j] = pi1[j];
            tau[j] = en * amu[j] * pi[j] - (en + 1.0) * pi0[j];
            s1a[j] += fn * (an * pi[j] + bn * tau[j]);
            s2a[j] += fn * (an * tau[j] + bn * pi[j]);
            s1b[j] += fn * p * (an * pi[j] - bn * tau[j]);
            s2b[j] += fn * p * (bn * pi[j] - an * tau[j]);
        }
        psi0 = psi1; psi1 = psi; chi0 = chi1; chi1 = chi;
        xi1 = cd(psi1, -chi1);
        for (int j = 0; j < nang; j++) {
            pi1[j] = ((2.0 * en + 1.0) * amu[j] * pi[j] - (en + 1.0) * pi0[j]) / en;
            pi0[j] = pi[j];
        }
    }
    // angles 0..90 from the forward sums, 90..180 from the mirrored ones (bhmie.py: concatenate(s_1, s_2[-2::-1]))
    for (int j = 0; j < nang; j++) {
        s1[2 * j] = s1a[j].real(); s1[2 * j + 1] = s1a[j].imag();
        s2[2 * j] = s2a[j].real(); s2[2 * j + 1] = s2a[j].imag();
    }
    for (int j = 0; j < nang - 1; j++) {
        const int src = nang - 2 - j, dst = nang + j;
        s1[2 * dst] = s1b[src].real(); s1[2 * dst + 1] = s1b[src].imag();
        s2[2 * dst] = s2b[src].real(); s2[2 * dst + 1] = s2b[src].imag();
    }
    if (q) {
        q[3] = 2.0 * gsca / qsca;
        q[1] = (2.0 / (x * x)) * qsca;
        q[0] = (4.0 / (x * x)) * s1[0];
        const double sb = std::hypot(s1[2 * (2 * nang - 2)], s1[2 * (2 * nang - 2) + 1]);
        q[2] = 4.0 * (sb / x) * (sb / x);
    }
    return 0;
}

}  // namespace ph
