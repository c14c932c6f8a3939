// Headless driver of the triangleparity handler (sms_b200/host/conshdlr_triangles.cpp) over the recording SCIP fake.
//   triangles_driver <instance.txt> <mode>
// instance.txt:  n m root nfixed sepacutpool / m lines "u v" in addEdge order / m lines x* by edge id / nfixed ids (unused)
// mode: sepalp | enfolp | sepasol | check | enfops | clone [infeasible_at_row]
// Prints "triangles <T>", one line per row SCIP received ("<pool|lp> <name> <rhs> <len> eid:coef ...") and "result <code>".
// Exit code 3: the handler could not be constructed (e.g. no CUDA device -- there is no CPU fallback).
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <stdexcept>
#include <string>

#include "sms/conshdlr/conshdlr_triangles.hpp"
#include "sms/probdata/mc_rooted.hpp"

int main(int argc, char **argv) {
    if (argc < 3) return 2;
    std::ifstream in(argv[1]);
    const std::string mode = argv[2];
    long n, m, root, nfixed, cutpool;
    in >> n >> m >> root >> nfixed >> cutpool;
    NetworKit::Graph g(n);
    for (long e = 0; e < m; ++e) {
        long u, v;
        in >> u >> v;
        g.addEdge(u, v);
    }
    g.indexEdges();
    Scip scip;
    ProbDataRootedMc prob(&scip, &g, root);
    scip.probdata = &prob;
    scip.lpsol.assign(m + n, 0.);
    for (long e = 0; e < m; ++e) in >> scip.lpsol[e];
    if (argc > 3) scip.infeasibleAtAddRow = std::atol(argv[3]);
    sms::ConshdlrTriangles *h = nullptr;
    try {
        h = new sms::ConshdlrTriangles(&scip, &g);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "handler construction failed: %s\n", e.what());
        return 3;
    }
    if (SCIPincludeObjConshdlr(&scip, h, TRUE) != SCIP_OKAY) return 4;
    *scip.boolParams.at("conshdlr/triangleparity/sepa_cutpool") = cutpool ? TRUE : FALSE;
    SCIP_CONSHDLR *ch = SCIPfindConshdlr(&scip, "triangleparity");
    if (!ch) return 5;
    SCIP_RESULT result = SCIP_DIDNOTRUN;
    SCIP_RETCODE rc = h->scip_initsol(&scip, ch, nullptr, 0);
    std::printf("triangles %lld\n", (long long)h->numTriangles());
    SCIP_Sol sol{scip.lpsol};
    if (rc == SCIP_OKAY) {
        if (mode == "sepalp")
            rc = h->scip_sepalp(&scip, ch, nullptr, 0, 0, &result);
        else if (mode == "enfolp")
            rc = h->scip_enfolp(&scip, ch, nullptr, 0, 0, FALSE, &result);
        else if (mode == "check")
            rc = h->scip_check(&scip, ch, nullptr, 0, &sol, TRUE, TRUE, FALSE, FALSE, &result);
        else if (mode == "sepasol" || mode == "enfops") {
            try {
                if (mode == "sepasol")
                    h->scip_sepasol(&scip, ch, nullptr, 0, 0, &sol, &result);
                else
                    h->scip_enfops(&scip, ch, nullptr, 0, 0, FALSE, FALSE, &result);
            } catch (const std::runtime_error &e) {
                std::printf("threw %s\n", e.what());
            }
        } else if (mode == "clone") {
            SCIP_Bool valid = FALSE;
            auto *c = h->clone(&scip, &valid);
            std::printf("clone %s valid %u\n", c ? "ok" : "null", valid);
            delete c;
        }
    }
    if (rc != SCIP_OKAY) {
        std::printf("retcode %d\n", (int)rc);
        return 6;
    }
    auto dump = [](const char *where, const std::vector<SCIP_ROW> &rows) {
        for (const auto &r : rows) {
            std::printf("%s %s %.17g %zu", where, r.name.c_str(), r.rhs, r.terms.size());
            for (const auto &t : r.terms) std::printf(" %d:%g", t.first->index, t.second);
            std::printf(" | lhs %g flags %u%u%u\n", r.lhs, r.local, r.modifiable, r.removable);
        }
    };
    dump("pool", scip.poolCuts);
    dump("lp", scip.lpRows);
    std::printf("result %d rows_created %ld rows_released %ld addvar %ld\n", (int)result, scip.nRowsCreated, scip.nRowsReleased, scip.nAddVar);
    delete h;
    return 0;
}
