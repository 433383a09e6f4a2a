// tests/binding/pfaai/algorithm_impl.hpp -- the reference-side binding of the B200 AJI path, as a
// drop-in for the reference's include/pfaai/algorithm_impl.hpp.
//
// Put this directory in front of the reference's include/ on the include path and every reference translation unit that says
//     #include "pfaai/algorithm_impl.hpp"        (src/main.cpp:33, tests/pfaai_tests.cpp:33)
// gets THIS ParFAAIImpl: the same class name, template parameters, constructor and public methods (computeJAC, computeAJI, run,
// getJAC, getAJI, genomePairToIndex; algorithm_impl.hpp:41-45,76-84,281-329 of the reference), with the work done by
// libpfaai_gpu.so through the C-ABI of include/pfaai_gpu.h.  The data-structure classes (ds_impl.hpp), the SQLite layer and the
// tests stay the reference's own, unmodified: oracle/Makefile compiles /root/reference/tests/pfaai_tests.cpp against this header
// into oracle/_ref/pfaai_tests_gpu.x and tests/test_gpu_reference_cases.py runs the reference's Catch2 cases through it.
//
// Only the interface (DataStructInterface, interface.hpp:225-327) is used, so one implementation serves the three drivers:
//   all-vs-all   ParFAAIData          qrySetSize() == tgtSetSize()
//   query subset ParFAAIQSubData      query ids = the genomes with isQryGenome(), in mapQueryId() order (= query-file order)
//   query DB     ParFAAIQryTgtData    refT() has qrySetSize() + tgtSetSize() columns (targets first); the host T is handed
//                                     over (the inner-joined F does not determine it) and PFAAI_FLAG_REF_COMPAT_QT reproduces the
//                                     labels initJAC gives the pairs (ds_impl.hpp:434-436)
// JAC and AJI come back in initJAC() order of the mode, which is what getJAC()/getAJI() callers index.
// No reference code is restated here: E is never built or read (refE() is ignored).
#pragma once
#include <cstdint>
#include <iostream>
#include <type_traits>
#include <vector>

#include "pfaai/interface.hpp"
#include "pfaai/utils.hpp"

#define PFAAI_GPU_REFERENCE_ERROR_CODES 1     // interface.hpp already defines PFAAI_OK ... PFAAI_ERR_CONSTRUCT
#include "pfaai_gpu.h"

template <typename IdType, typename ValueType, typename DSIT = DefaultDataStructInterface<IdType>>
class ParFAAIImpl {
  public:
    using IdPairType = typename DSIT::IdPairType;
    using IdMatrixType = typename DSIT::IdMatrixType;
    using JACType = typename DSIT::JACType;
    using DSIfx = DSIT;

  private:
    static_assert(std::is_same<IdType, int>::value && std::is_same<ValueType, double>::value,
                  "the GPU path computes int32 counts and fp64 sums, as the reference's IdType = int / ValueType = double build");
    static_assert(sizeof(IdPairType) == sizeof(pfaai_pg_t), "DPair<int,int> and pfaai_pg_t are both two int32 (protein, genome)");

    const DSIfx& c_pfDSRef;
    pfaai_gpu_ctx* m_ctx = nullptr;
    int m_mode = PFAAI_ALL_VS_ALL;
    unsigned m_flags = 0;
    std::vector<int32_t> m_qryIds;
    bool m_done = false;
    std::vector<JACType> m_JAC;
    std::vector<ValueType> m_AJI, m_AJIResult, m_matrix;

    int fail(const char* where) {
        std::cerr << "ParFAAIImpl (GPU binding): " << where << ": " << (m_ctx ? pfaai_gpu_last_error(m_ctx) : "no context") << std::endl;
        return PFAAI_ERR_CONSTRUCT;
    }

    int computeAll() {
        if (m_done) return PFAAI_OK;
        const std::vector<IdType>& Lc = c_pfDSRef.refLc();
        const std::vector<IdPairType>& F = c_pfDSRef.refF();
        const IdMatrixType& T = c_pfDSRef.refT();
        const int nQ = c_pfDSRef.qrySetSize(), nT = c_pfDSRef.tgtSetSize();
        const int P = static_cast<int>(T.rows()), G = static_cast<int>(T.cols());
        const int32_t* hostT = nullptr;
        if (G == nQ + nT) {                        // query DB against target DB: the query genomes follow the targets
            m_mode = PFAAI_QUERY_DB;
            m_flags = PFAAI_FLAG_REF_COMPAT_QT;
            hostT = T.data().data();
            for (int q = 0; q < nQ; q++) m_qryIds.push_back(nT + q);
        } else if (nQ == nT) {
            m_mode = PFAAI_ALL_VS_ALL;
        } else {                                   // a subset of the database's own genomes, in query-file order
            m_mode = PFAAI_QUERY_SUBSET;
            m_qryIds.assign(nQ, -1);
            for (int g = 0; g < G; g++)
                if (c_pfDSRef.isQryGenome(g)) m_qryIds[c_pfDSRef.mapQueryId(g)] = g;
        }
        if (pfaai_gpu_create(&m_ctx, /*device=*/0, /*flags=*/0) != PFAAI_OK) return fail("pfaai_gpu_create");
        if (pfaai_gpu_upload_index(m_ctx, P, G, reinterpret_cast<const pfaai_pg_t*>(F.data()), Lc.data(),
                                   static_cast<int64_t>(F.size()), hostT) != PFAAI_OK)
            return fail("pfaai_gpu_upload_index");
        if (pfaai_gpu_build(m_ctx) != PFAAI_OK) return fail("pfaai_gpu_build");
        const int64_t nPairs = pfaai_gpu_n_pairs(m_mode, G, nQ);
        if (nPairs != static_cast<int64_t>(c_pfDSRef.nGenomePairs())) return fail("pair count differs from nGenomePairs()");
        std::vector<pfaai_jac_t> jac(nPairs);
        m_AJIResult.assign(nPairs, 0.0);
        // shape of printOutput's matrix per mode (pfaai_gpu_matrix_shape reports the same after the compute)
        const int64_t rows = (m_mode == PFAAI_ALL_VS_ALL) ? G : nQ;
        const int64_t cols = (m_mode == PFAAI_QUERY_DB) ? nT : G;
        m_matrix.assign(static_cast<size_t>(rows) * cols, 0.0);
        if (pfaai_gpu_compute(m_ctx, m_mode, m_qryIds.empty() ? nullptr : m_qryIds.data(), static_cast<int>(m_qryIds.size()), m_flags,
                              m_matrix.data(), jac.data(), m_AJIResult.data(), nullptr) != PFAAI_OK)
            return fail("pfaai_gpu_compute");
        m_JAC.resize(nPairs);
        for (int64_t i = 0; i < nPairs; i++) {
            m_JAC[i].genomeA = jac[i].gA;  m_JAC[i].genomeB = jac[i].gB;
            m_JAC[i].S = jac[i].S;         m_JAC[i].N = jac[i].N;
        }
        m_done = true;
        return PFAAI_OK;
    }

  public:
    explicit ParFAAIImpl(const DSIfx& fDataRef) : c_pfDSRef(fDataRef) {}
    ParFAAIImpl(const ParFAAIImpl&) = delete;
    ParFAAIImpl& operator=(const ParFAAIImpl&) = delete;
    ~ParFAAIImpl() { if (m_ctx) pfaai_gpu_destroy(m_ctx); }

    const std::vector<JACTuple<IdType>>& getJAC() const { return m_JAC; }
    const std::vector<ValueType>& getAJI() const { return m_AJI; }
    // the dense output of printOutput (src/main.cpp:133-175), row = mapQueryId, column = mapTargetId, ready to be written
    const std::vector<ValueType>& getAJIMatrix() const { return m_matrix; }

    inline IdType genomePairToIndex(IdType genomeA, IdType genomeB) const { return c_pfDSRef.genomePairToIndex(genomeA, genomeB); }

    // one launch sequence produces JAC, AJI and the matrix; the two phase calls keep the reference's observable order:
    // getAJI() is empty until computeAJI() has been called
    int computeJAC() { return computeAll(); }
    int computeAJI() {
        const int rc = computeAll();
        if (rc == PFAAI_OK) m_AJI = m_AJIResult;
        return rc;
    }
    int run() {
        computeJAC();
        return computeAJI();
    }
};
