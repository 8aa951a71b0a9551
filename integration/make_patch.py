#!/usr/bin/env python
"""Regenerates integration/iqtree-1.4.3-lkcuda.patch from the edits below.

    python integration/make_patch.py [/root/reference]

The edits are anchored on short unique strings of the IQ-TREE 1.4.3 sources (line endings preserved: phylotree.h / phylotree.cpp are
CRLF files).  The result is a plain unified diff a maintainer applies with `patch --binary -p1` in the IQ-TREE source directory; together with
integration/phylotree_cuda.cpp (added to the source list) and include/lkcuda.h it is the whole reference-side binding.
"""
import difflib
import os
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

EDITS = {
    "tools.h": [
        ("    bool lk_no_avx;\n",
         "    bool lk_no_avx;\n"
         "\n"
         "    /** TRUE to run the likelihood kernels on the GPU through lkcuda.h (option -cuda), default: FALSE */\n"
         "    bool lk_cuda;\n"
         "\n"
         "    /** CUDA device ordinal for -cuda, default: 0 */\n"
         "    int cuda_device;\n"),
    ],
    "tools.cpp": [
        ("    params.lk_no_avx = false;\n",
         "    params.lk_no_avx = false;\n    params.lk_cuda = false;\n    params.cuda_device = 0;\n"),
        ("\t\t\tif (strcmp(argv[cnt], \"-noavx\") == 0) {\n\t\t\t\tparams.lk_no_avx = true;\n\t\t\t\tcontinue;\n\t\t\t}\n",
         "\t\t\tif (strcmp(argv[cnt], \"-noavx\") == 0) {\n\t\t\t\tparams.lk_no_avx = true;\n\t\t\t\tcontinue;\n\t\t\t}\n"
         "\t\t\tif (strcmp(argv[cnt], \"-cuda\") == 0) {\n"
         "\t\t\t\tparams.lk_cuda = true;\n"
         "\t\t\t\tif (cnt + 1 < argc && isdigit(argv[cnt+1][0]))\n"
         "\t\t\t\t\tparams.cuda_device = convert_int(argv[++cnt]);\n"
         "\t\t\t\tcontinue;\n"
         "\t\t\t}\n"),
    ],
    # the reference's dispatcher keeps its body under a new name; the new setLikelihoodKernel (phylotree_cuda.cpp) calls it first
    "phylotreesse.cpp": [
        ("void PhyloTree::setLikelihoodKernel(LikelihoodKernel lk) {\n    setParsimonyKernel(lk);",
         "void PhyloTree::setLikelihoodKernelCPU(LikelihoodKernel lk) {\n    setParsimonyKernel(lk);"),
    ],
    "phylotree.h": [
        ("    /** get the number of bytes occupied by scale_num */\n    int getScaleNumBytes();\n",
         "    /** get the number of bytes occupied by scale_num */\n    int getScaleNumBytes();\n"
         "\n"
         "    /****************************************************************************\n"
         "            B200 likelihood engine behind the kernel pointers (phylotree_cuda.cpp, lkcuda.h; option -cuda)\n"
         "     ****************************************************************************/\n"
         "    struct LkCuda;\n"
         "    /** engine, device arena and recorded nodes of this tree; NULL while the buffers are host buffers */\n"
         "    LkCuda *lk_cuda;\n"
         "    /** the CPU dispatcher (the original setLikelihoodKernel) */\n"
         "    void setLikelihoodKernelCPU(LikelihoodKernel lk);\n"
         "    bool lkCudaWanted();\n"
         "    void lkCudaPrepareArena();\n"
         "    void lkCudaAfterInit();\n"
         "    void lkCudaRelease();\n"
         "    int lkCudaVector(PhyloNeighbor *nei);\n"
         "    int lkCudaScaleNumBytes();\n"
         "    void lkCudaSyncInputs();\n"
         "    void lkCudaFlush();\n"
         "    void lkCudaBranchEnds(PhyloNeighbor *&dad_branch, PhyloNode *&dad, int &dad_end, int &node_vec);\n"
         "    void computePartialLikelihoodCuda(PhyloNeighbor *dad_branch, PhyloNode *dad);\n"
         "    double computeLikelihoodBranchCuda(PhyloNeighbor *dad_branch, PhyloNode *dad);\n"
         "    void computeLikelihoodDervCuda(PhyloNeighbor *dad_branch, PhyloNode *dad, double &df, double &ddf);\n"
         "    double computeLikelihoodFromBufferCuda();\n"
         "    double lkCudaScaleFactor(PhyloNeighbor *nei);\n"
         "    void lkCudaFetchVector(PhyloNeighbor *nei, double *partial_lh, UBYTE *scale_num);\n"
         "    void lkCudaFetchTheta(double *theta);\n"
         "    void lkCudaSyncPatternLh();\n"
         "    void lkCudaPatternLh(double *pattern_lh);\n"
         "    double computePatternLhCatCuda(SiteLoglType wsl);\n"
         "    void computePatternLikelihoodCuda(double *ptn_lh, double *cur_logl, double *ptn_lh_cat, SiteLoglType wsl);\n"
         "    void lkCudaStats(uint64_t &flushes, uint64_t &ops, uint64_t &launches);\n"
         "    void lkCudaProfile(bool on, bool reset);\n"
         "    bool lkCudaKernelStats(int kind, uint64_t &launches, uint64_t &bytes, uint64_t &flops, double &device_ms);\n"),
    ],
    "phylotree.cpp": [
        ("    var_matrix = NULL;\n    setLikelihoodKernel(LK_EIGEN_SSE);",
         "    var_matrix = NULL;\n    lk_cuda = NULL;\n    setLikelihoodKernel(LK_EIGEN_SSE);"),
        ("PhyloTree::~PhyloTree() {\n",
         "PhyloTree::~PhyloTree() {\n    lkCudaRelease();\n"),
        ("    double wall_start_time = getRealTime();\n    initializeAllPartialLh(index, indexlh);\n",
         "    double wall_start_time = getRealTime();\n    lkCudaPrepareArena();\n    initializeAllPartialLh(index, indexlh);\n    lkCudaAfterInit();\n"),
        ("void PhyloTree::deleteAllPartialLh() {\n\n",
         "void PhyloTree::deleteAllPartialLh() {\n\n\tlkCudaRelease();\n"),
        ("int PhyloTree::getScaleNumBytes() {\n",
         "int PhyloTree::getScaleNumBytes() {\n\tif (lk_cuda) // vectors of the device arena are scale_block_size apart, the NNI vectors included\n\t\treturn lkCudaScaleNumBytes();\n"),
        ("    if (pattern_lh)\n        memmove(pattern_lh, _pattern_lh, aln->size() * sizeof(double));\n",
         "    if (pattern_lh && lk_cuda) {\n        lkCudaPatternLh(pattern_lh);\n        return score;\n    }\n"
         "    if (pattern_lh)\n        memmove(pattern_lh, _pattern_lh, aln->size() * sizeof(double));\n"),
        ("//    } else \n    if (!getModel()->isMixture())\n",
         "//    } else \n    if (lk_cuda)\n        return computePatternLhCatCuda(wsl);\n    if (!getModel()->isMixture())\n"),
        ("    int nptn = aln->getNPattern();\n    int i;\n    int ncat = getNumLhCat(wsl);\n    if (ptn_lh_cat) {\n",
         "    if (lk_cuda) {\n        computePatternLikelihoodCuda(ptn_lh, cur_logl, ptn_lh_cat, wsl);\n        return;\n    }\n"
         "    int nptn = aln->getNPattern();\n    int i;\n    int ncat = getNumLhCat(wsl);\n    if (ptn_lh_cat) {\n"),
    ],
}


def patched(name):
    raw = open(os.path.join(REF, name), "rb").read().decode("latin-1")
    crlf = "\r\n" in raw
    text = raw.replace("\r\n", "\n")
    for old, new in EDITS[name]:
        if text.count(old) != 1:
            raise SystemExit("%s: anchor found %d times: %r" % (name, text.count(old), old[:60]))
        text = text.replace(old, new)
    if crlf:
        text = text.replace("\n", "\r\n")
    return raw, text


def main():
    global REF
    if len(sys.argv) > 1:
        REF = sys.argv[1]
    out = []
    for name in sorted(EDITS):
        a, b = patched(name)
        diff = difflib.unified_diff(a.splitlines(True), b.splitlines(True), "a/" + name, "b/" + name, n=3)
        out.extend(diff)
    data = "".join(out).encode("latin-1")
    dst = os.path.join(HERE, "iqtree-1.4.3-lkcuda.patch")
    open(dst, "wb").write(data)
    print("%s: %d bytes, %d files" % (dst, len(data), len(EDITS)))


if __name__ == "__main__":
    main()
