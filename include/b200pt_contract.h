/* b200pt_contract.h — the one piece of arithmetic that the GLSL text of the reference leaves to the implementation and
 * that the CUDA path, the CPU oracle and the reference shaders compiled as C++ (oracle/ref_glsl/glsl_compat.h) must all
 * evaluate identically: sin() and cos() of an fp32 argument (brdf.glsl:54,58,108,112: phi = 2*pi*xi).
 *
 * GLSL gives sin/cos no exact definition (the Vulkan precision rules only bound the absolute error by 2^-11 on
 * [-pi, pi]), so parity needs a definition that is *computable bit for bit on both machines*: a library call is not
 * (glibc's and CUDA's double sin() agree only to the last place, and after rounding to fp32 about one call in 2^29
 * differs).  This header is that definition: argument reduction by multiples of pi/2 (two-constant Cody-Waite, exact
 * for |k| < 2^20) and the fdlibm kernel polynomials, written with explicit IEEE double +, * in a fixed order — no
 * library call, no FMA (both sides compile with contraction off).  The double result is within 1 ulp(double) of the true
 * value, i.e. the rounded float is the correctly rounded one except within ~2^-29 of a rounding boundary; what matters
 * here is that every side gets the SAME float.  Valid for |x| < 2^20 (the shaders use [0, 2*pi]). */
#ifndef B200PT_CONTRACT_H
#define B200PT_CONTRACT_H

#ifdef __CUDACC__
#define PT_CONTRACT_FN __host__ __device__ inline
#else
#define PT_CONTRACT_FN inline
#endif

PT_CONTRACT_FN void pt_contract_sincos(float x, float* s_out, float* c_out) {
    const double xd = (double)x;
    /* k = nearest integer to x / (pi/2): adding and subtracting 1.5 * 2^52 rounds to nearest-even in IEEE arithmetic */
    const double big = 6755399441055744.0;
    const double kd = (xd * 0.63661977236758134308 + big) - big;
    /* pi/2 = hi + lo with hi carrying 33 bits: kd * hi is exact for |kd| < 2^20 */
    const double r = (xd - kd * 1.57079632673412561417e+00) - kd * 6.07710050650619224932e-11;
    const double z = r * r;
    /* fdlibm __kernel_sin / __kernel_cos on [-pi/4, pi/4] */
    const double ps = -1.66666666666666324348e-01 + z * (8.33333333332248946124e-03 + z * (-1.98412698298579493134e-04 +
                      z * (2.75573137070700676789e-06 + z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10))));
    const double pc = 4.16666666666666019037e-02 + z * (-1.38888888888741095749e-03 + z * (2.48015872894767294178e-05 +
                      z * (-2.75573143513906633035e-07 + z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11))));
    const double sr = r + r * (z * ps);
    const double cr = (1.0 - 0.5 * z) + (z * z) * pc;
    const long long k = (long long)kd;
    const int q = (int)(k & 3);
    const double sd = (q == 0) ? sr : (q == 1) ? cr : (q == 2) ? -sr : -cr;
    const double cd = (q == 0) ? cr : (q == 1) ? -sr : (q == 2) ? -cr : sr;
    *s_out = (float)sd;
    *c_out = (float)cd;
}

#endif
