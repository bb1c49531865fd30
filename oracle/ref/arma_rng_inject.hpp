// TEST INFRASTRUCTURE ONLY (oracle/): Armadillo's documented RNG replacement hook.
//
// Compiled in with -DARMA_RNG_ALT=arma_rng_inject.hpp (armadillo-8.400.0/include/armadillo:122-123;
// dispatch at armadillo_bits/arma_rng.hpp:41-55,324-360).  Two sources of standard normals:
//   * INJECT: pop FP64 draws, in consumption order, from a caller-supplied buffer -- this is the
//     "noise draws injected from a file" mode of BASELINE.json; the reference then consumes
//     3N (initParticles, MCSimulator.h:290) + (3 + 8) per step (:754, :786-789) values.
//   * MT: std::mt19937_64 + std::normal_distribution<double>, the engine Armadillo itself uses
//     when built with its wrapper (armadillo_bits/arma_rng_cxx11.hpp) -- used for the run-to-run
//     distribution of the reference estimator (99 % CI for Philox mode).
#ifndef ORACLE_ARMA_RNG_INJECT_HPP
#define ORACLE_ARMA_RNG_INJECT_HPP

#include <random>
#include <cstdlib>
#include <cstdio>

struct oracle_rng_state {
    const double* draws;
    size_t n_draws;
    size_t pos;
    bool inject;
    bool underflow;
    std::mt19937_64 engine;
    std::normal_distribution<double> normal;
    std::uniform_real_distribution<double> uniform;
    oracle_rng_state() : draws(nullptr), n_draws(0), pos(0), inject(false), underflow(false) {}
};

inline oracle_rng_state& oracle_rng() {
    static oracle_rng_state s;
    return s;
}

class arma_rng_alt {
public:
    typedef unsigned long long seed_type;

    // The reference reseeds from the clock in its constructor (MCSimulator.h:141); the oracle keeps
    // its own seed so runs are repeatable.
    inline static void set_seed(const seed_type) {}

    inline static int randi_val() { return int(oracle_rng().engine() & 0x7fffffff); }
    inline static double randu_val() { return oracle_rng().uniform(oracle_rng().engine); }

    inline static double randn_val() {
        oracle_rng_state& s = oracle_rng();
        if (s.inject) {
            if (s.pos >= s.n_draws) { s.underflow = true; return 0.0; }
            return s.draws[s.pos++];
        }
        return s.normal(s.engine);
    }

    template <typename eT>
    inline static void randn_dual_val(eT& out1, eT& out2) {
        out1 = eT(randn_val());
        out2 = eT(randn_val());
    }

    template <typename eT>
    inline static void randi_fill(eT* mem, const unsigned long long N, const int a, const int b) {
        std::uniform_int_distribution<int> d(a, b);
        for (unsigned long long i = 0; i < N; ++i) mem[i] = eT(d(oracle_rng().engine));
    }

    inline static int randi_max_val() { return 0x7fffffff; }
};

#endif
