// oracle/guac_oracle_assembly.hpp — TEST INFRASTRUCTURE ONLY (see guac_oracle.hpp).
// CPU restatement of the germline-assembly path (SURVEY.md §8 rows A17–A20):
//   M/assembly/DeBruijnGraph.scala, M/assembly/AssemblyUtils.scala:50-174,
//   M/alignment/ReadAlignment.scala + AlignmentState.scala,
//   M/commands/GermlineAssemblyCaller.scala:103-219.
// Pinned by DeBruijnGraphSuite / ReadAlignmentSuite vectors (tests/test_oracle_assembly.py).  The end-to-end
// GermlineAssemblyCallerSuite needs BAM / FASTA fixtures that are absent from the checkout
// (.MISSING_LARGE_BLOBS): the caller closure is parity-UNPINNED beyond its components.
//
// Iteration orders that come out of Scala hash collections in the reference (children / parents lists,
// `allNodes.head`) are replaced by lexicographic k-mer order here and in the CUDA path; path SETS do not
// depend on them unless the 17-path cap bites.
#pragma once
#include <map>
#include <set>

#include "guac_oracle.hpp"

namespace guac_oracle {

// ---------------------------------------------------------------------------------
// DeBruijnGraph (M/assembly/DeBruijnGraph.scala:10-464)
// ---------------------------------------------------------------------------------
struct DeBruijnGraph {
  int kmerSize;
  std::map<std::string, int> kmerCounts;
  std::map<std::string, std::vector<std::string>> prefixTable, suffixTable;  // :18-35
  std::map<std::string, std::pair<std::string, int>> mergeIndex;              // :37

  std::string kmerPrefix(const std::string& s) const { return s.substr(0, std::min<size_t>(s.size(), kmerSize - 1)); }  // :43-45
  std::string kmerSuffix(const std::string& s) const {                                                                   // :48-50
    const size_t n = std::min<size_t>(s.size(), kmerSize - 1);
    return s.substr(s.size() - n);
  }
  DeBruijnGraph(int k, std::map<std::string, int> counts) : kmerSize(k), kmerCounts(std::move(counts)) {
    for (auto& kv : kmerCounts) {
      prefixTable[kmerPrefix(kv.first)].push_back(kv.first);
      suffixTable[kmerSuffix(kv.first)].push_back(kv.first);
    }
  }
  std::vector<std::string> children(const std::string& node) const {  // :322-326
    auto it = prefixTable.find(kmerSuffix(node));
    return it == prefixTable.end() ? std::vector<std::string>{} : it->second;
  }
  std::vector<std::string> parents(const std::string& node) const {  // :330-334
    auto it = suffixTable.find(kmerPrefix(node));
    return it == suffixTable.end() ? std::vector<std::string>{} : it->second;
  }
  void removeKmer(const std::string& kmer) {  // :58-77
    kmerCounts.erase(kmer);
    auto rm = [&](std::map<std::string, std::vector<std::string>>& table, const std::string& key) {
      auto it = table.find(key);
      if (it == table.end()) return;
      auto& v = it->second;
      v.erase(std::remove(v.begin(), v.end(), kmer), v.end());
      if (v.empty()) table.erase(it);
    };
    rm(prefixTable, kmerPrefix(kmer));
    rm(suffixTable, kmerSuffix(kmer));
  }
  void pruneKmers(int minSupport) {  // :82-86
    std::vector<std::string> drop;
    for (auto& kv : kmerCounts)
      if (kv.second < minSupport) drop.push_back(kv.first);
    for (auto& k : drop) removeKmer(k);
  }
  // findMergeable (:136-168); returns the list most-recent-first like the Scala
  std::vector<std::string> findMergeable(const std::string& kmer, bool searchForward, bool avoidLoops = true) const {
    auto nextFunc = [&](const std::string& n) { return searchForward ? children(n) : parents(n); };
    auto prevFunc = [&](const std::string& n) { return searchForward ? parents(n) : children(n); };
    std::string current = kmer;
    std::set<std::string> visited{current};
    auto nextNodes = [&](const std::string& n) {
      auto v = nextFunc(n);
      if (avoidLoops) v.erase(std::remove_if(v.begin(), v.end(), [&](const std::string& x) { return visited.count(x) > 0; }), v.end());
      return v;
    };
    auto next = nextNodes(current);
    std::vector<std::string> mergeable{current};
    while (next.size() == 1 && prevFunc(next[0]).size() == 1) {
      current = next[0];
      visited.insert(current);
      mergeable.insert(mergeable.begin(), current);
      next = nextNodes(current);
    }
    return mergeable;
  }
  std::vector<std::string> mergeForward(const std::string& k) const {  // :170
    auto v = findMergeable(k, true);
    std::reverse(v.begin(), v.end());
    return v;
  }
  std::vector<std::string> mergeBackward(const std::string& k) const { return findMergeable(k, false); }  // :171

  static std::string mergeOverlappingSequences(const std::vector<std::string>& sequences, int overlapSize) {  // :384-388
    if (sequences.empty()) return "";
    std::string out = sequences[0];
    for (size_t i = 1; i < sequences.size(); ++i) {
      const std::string& s = sequences[i];
      const int take = (int)s.size() - overlapSize + 1;
      if (take > 0) out += s.substr(s.size() - take);
    }
    return out;
  }

  void mergeNodes() {  // :91-126
    std::set<std::string> allNodes;
    for (auto& kv : kmerCounts) allNodes.insert(kv.first);
    while (!allNodes.empty()) {
      const std::string node = *allNodes.begin();
      auto forwardUniquePath = mergeForward(node);
      auto backwardUniquePath = mergeBackward(node);
      std::vector<std::string> full = backwardUniquePath;
      full.insert(full.end(), forwardUniquePath.begin() + 1, forwardUniquePath.end());
      if (full.size() > 1) {
        for (auto& k : full) {
          allNodes.erase(k);
          if (kmerCounts.count(k)) removeKmer(k);
        }
        const std::string merged = mergeOverlappingSequences(full, kmerSize);
        for (size_t i = 0; i < full.size(); ++i) mergeIndex[full[i]] = {merged, (int)i};
        auto& pl = prefixTable[kmerPrefix(merged)];
        pl.insert(pl.begin(), merged);
        auto& sl = suffixTable[kmerSuffix(merged)];
        sl.insert(sl.begin(), merged);
        kmerCounts[merged] = (kmerCounts.count(merged) ? kmerCounts[merged] : 0) + 1;
      } else {
        allNodes.erase(node);
      }
    }
  }

  using Path = std::vector<std::string>;
  // depthFirstSearch (:183-300); paths are returned most-recent-first like the Scala list
  std::vector<Path> depthFirstSearch(const std::string& source, const std::string& sink, int minPathLength = 1,
                                     int maxPathLength = std::numeric_limits<int>::max(), int maxPaths = 10,
                                     bool avoidLoops = true) const {
    if ((int)source.size() != kmerSize) throw InvalidInput("Source kmer has size != kmerSize");
    if ((int)sink.size() != kmerSize) throw InvalidInput("Sink kmer has size != kmerSize");
    std::vector<Path> paths;
    std::set<std::string> visited;
    std::vector<std::string> frontier;  // stack: back() is the top
    auto srcIt = mergeIndex.find(source);
    auto sinkIt = mergeIndex.find(sink);
    if (srcIt != mergeIndex.end()) {
      const std::string& mergedNode = srcIt->second.first;
      const int pathIndex = srcIt->second.second;
      if (sinkIt != mergeIndex.end() && sinkIt->second.first == mergedNode && sinkIt->second.second > pathIndex)
        return {Path{mergedNode.substr(pathIndex, sinkIt->second.second + kmerSize - pathIndex)}};
      visited.insert(mergedNode);
      frontier.push_back(mergedNode.substr(pathIndex));
    } else {
      frontier.push_back(source);
    }
    Path currentPath;  // front() is the most recent node (Scala list head)
    std::vector<int> lastBranchIndex, lastBranchChildren;
    int numBacktracks = 0;
    while (!frontier.empty() && (int)paths.size() < maxPaths) {
      const std::string next = frontier.back();
      frontier.pop_back();
      currentPath.insert(currentPath.begin(), next);
      const bool foundMergedSink = sinkIt != mergeIndex.end() && sinkIt->second.first == next;
      const bool foundSink = next == sink || foundMergedSink;
      auto nextNodes = children(next);
      if (avoidLoops) nextNodes.erase(std::remove_if(nextNodes.begin(), nextNodes.end(), [&](const std::string& x) { return visited.count(x) > 0; }), nextNodes.end());
      if (!foundSink && !nextNodes.empty() && (int)currentPath.size() < maxPathLength) {
        if (nextNodes.size() > 1) {
          lastBranchIndex.push_back((int)currentPath.size());
          lastBranchChildren.push_back((int)nextNodes.size());
        }
        for (auto& n : nextNodes) frontier.push_back(n);  // pushAll: the last child ends on top
        visited.insert(next);
      } else {
        if (foundSink && (int)currentPath.size() >= minPathLength) {
          if (foundMergedSink) {
            const std::string& mergedNode = sinkIt->second.first;
            const int mergedPathEndIdx = sinkIt->second.second + kmerSize;
            const int drop = (int)mergedNode.size() - mergedPathEndIdx;
            currentPath[0] = currentPath[0].substr(0, std::max(0, (int)currentPath[0].size() - drop));
          }
          Path p(currentPath.rbegin(), currentPath.rend());
          paths.insert(paths.begin(), p);
        }
        if (!lastBranchIndex.empty()) {
          const int backtrackIndex = lastBranchIndex.back();
          const int n = (int)currentPath.size() - backtrackIndex;
          for (int i = 0; i < n; ++i) visited.erase(currentPath[i]);
          currentPath.erase(currentPath.begin(), currentPath.begin() + std::max(0, n));
          numBacktracks += 1;
          if (numBacktracks == lastBranchChildren.back() - 1) {
            lastBranchIndex.pop_back();
            lastBranchChildren.pop_back();
            numBacktracks = 0;
          }
        }
      }
    }
    return paths;
  }
};

// breeze mean over Seq[Float] (running mean in single precision)
inline float breezeMeanFloat(const uint8_t* q, int n) {
  float mu = 0;
  long cnt = 0;
  for (int i = 0; i < n; ++i) {
    cnt += 1;
    const float d = (float)q[i] - mu;
    mu = mu + 1.0f / cnt * d;
  }
  return mu;
}

// object DeBruijnGraph.apply (:342-373): k-mer counts over the reads, mean-quality gate, minOccurrence
inline DeBruijnGraph buildGraph(const std::vector<const MappedRead*>& reads, int kmerSize, int minOccurrence, int minMeanKmerBaseQuality,
                                bool mergeNodes) {
  std::map<std::string, int> counts;
  for (auto* read : reads) {
    const int L = (int)read->sequence.size();
    if (L == 0) continue;
    const int nWindows = L < kmerSize ? 1 : L - kmerSize + 1;  // Iterator.sliding yields one short group when L < k
    for (int index = 0; index < nWindows; ++index) {
      const int len = std::min(kmerSize, L - index);
      const float meanBaseQuality = breezeMeanFloat(read->baseQualities.data() + index, len);
      if (meanBaseQuality > minMeanKmerBaseQuality) counts[read->sequence.substr(index, len)] += 1;
    }
  }
  std::map<std::string, int> kept;
  for (auto& kv : counts)
    if (kv.second >= minOccurrence) kept.insert(kv);
  DeBruijnGraph g(kmerSize, kept);
  if (mergeNodes) g.mergeNodes();
  return g;
}

// discoverPathsFromReads (:437-464) followed by mergeOverlappingSequences per path (AssemblyUtils.scala:84)
inline std::vector<std::string> discoverPathSequences(const std::vector<const MappedRead*>& reads, const std::string& referenceSequence,
                                                      int kmerSize, int minOccurrence, int maxPaths, int minMeanKmerBaseQuality) {
  if ((int)referenceSequence.size() < kmerSize) throw InvalidInput("reference window shorter than kmerSize");
  DeBruijnGraph g = buildGraph(reads, kmerSize, minOccurrence, minMeanKmerBaseQuality, true);
  auto paths = g.depthFirstSearch(referenceSequence.substr(0, kmerSize), referenceSequence.substr(referenceSequence.size() - kmerSize), 1,
                                  std::numeric_limits<int>::max(), maxPaths);
  std::vector<std::string> out;
  for (auto& p : paths) out.push_back(DeBruijnGraph::mergeOverlappingSequences(p, kmerSize));
  return out;
}

// ---------------------------------------------------------------------------------
// ReadAlignment (M/alignment/ReadAlignment.scala:14-195), AlignmentState (M/alignment/AlignmentState.scala)
// ---------------------------------------------------------------------------------
enum AlignmentState : uint8_t { ST_MATCH = 0, ST_MISMATCH = 1, ST_INSERTION = 2, ST_DELETION = 3, ST_NONE = 255 };
inline bool isGapAlignment(uint8_t s) { return s == ST_INSERTION || s == ST_DELETION; }

struct ReadAlignmentResult {
  std::vector<uint8_t> alignments;  // states from the start of the sequence
  int refStartIdx = 0, refEndIdx = 0, alignmentScore = 0;
  double rawScore = 0;
  std::string toCigarString() const {  // :22-43
    if (alignments.empty()) return "";
    std::string out;
    const char* keys = "=XID";
    int run = 1;
    for (size_t i = 1; i <= alignments.size(); ++i) {
      if (i < alignments.size() && alignments[i] == alignments[i - 1]) {
        ++run;
      } else {
        out += std::to_string(run) + keys[alignments[i - 1]];
        run = 1;
      }
    }
    return out;
  }
};

// scoreAlignmentPaths (:92-195) with back-pointers instead of a List per cell; apply (:64-90)
inline ReadAlignmentResult readAlignment(const std::string& sequence, const std::string& reference, double mismatchProbability = std::exp(-4.0),
                                         double openGapProbability = std::exp(-6.0), double closeGapProbability = 1 - std::exp(-1.0)) {
  const double logMismatchPenalty = -std::log(mismatchProbability);
  const double logOpenGapPenalty = -std::log(openGapProbability);
  const double noGapPenalty = -std::log(1 - openGapProbability);
  const double logCloseGapPenalty = -std::log(closeGapProbability);
  const double logContinueGapPenalty = -std::log(1 - closeGapProbability);
  const int S = (int)sequence.size(), Rn = (int)reference.size();
  // cell (s, r): best path for sequence[s..] starting at reference index r; row S is the initial state
  std::vector<double> score((size_t)(S + 1) * (Rn + 1), 0.0);
  std::vector<int> endIdx((size_t)(S + 1) * (Rn + 1), 0);
  std::vector<uint8_t> head((size_t)(S + 1) * (Rn + 1), ST_NONE);  // state taken at this cell = head of its path list
  auto at = [&](int s, int r) { return (size_t)s * (Rn + 1) + r; };
  for (int r = 0; r <= Rn; ++r) endIdx[at(S, r)] = r;
  auto transitionPenalty = [&](uint8_t nextState, uint8_t prev, bool isEndState) {  // :118-130
    const bool hasPrev = prev != ST_NONE;
    const bool openGap = !(hasPrev && prev == nextState) && isGapAlignment(nextState);
    const bool closeGap = hasPrev && nextState != prev && isGapAlignment(prev);
    const bool continueGap = hasPrev && prev == nextState && isGapAlignment(nextState);
    const bool mismatch = nextState == ST_MISMATCH;
    return (openGap ? logOpenGapPenalty : 0) + (closeGap ? logCloseGapPenalty : 0) +
           (continueGap ? logContinueGapPenalty : (mismatch ? noGapPenalty + logMismatchPenalty : noGapPenalty)) +
           ((isEndState && isGapAlignment(nextState)) ? logCloseGapPenalty : 0);
  };
  for (int s = S - 1; s >= 0; --s)
    for (int r = Rn; r >= 0; --r) {
      const bool isEndState = s == 0;
      bool have = false;
      double best = 0;
      uint8_t bestState = ST_NONE;
      int bestEnd = 0;
      auto consider = [&](uint8_t st, size_t from) {
        const double sc = score[from] + transitionPenalty(st, head[from], isEndState);
        if (!have || sc < best) {  // minBy keeps the first of equal minima
          have = true;
          best = sc;
          bestState = st;
          bestEnd = endIdx[from];
        }
      };
      // candidate order of the Scala: (s+1, r) -> Insertion, (s, r+1) -> Deletion, (s+1, r+1) -> Match/Mismatch
      consider(ST_INSERTION, at(s + 1, r));
      if (r + 1 <= Rn) consider(ST_DELETION, at(s, r + 1));
      if (r + 1 <= Rn) consider(sequence[s] != reference[r] ? ST_MISMATCH : ST_MATCH, at(s + 1, r + 1));
      score[at(s, r)] = best;
      head[at(s, r)] = bestState;
      endIdx[at(s, r)] = bestEnd;
    }
  ReadAlignmentResult res;
  int bestR = 0;
  for (int r = 1; r <= Rn; ++r)
    if (score[at(0, r)] < score[at(0, bestR)]) bestR = r;
  res.refStartIdx = bestR;
  res.refEndIdx = endIdx[at(0, bestR)];
  res.rawScore = score[at(0, bestR)];
  res.alignmentScore = (int)res.rawScore;  // score.toInt
  // walk the back-pointers
  int s = 0, r = bestR;
  while (s < S || (head[at(s, r)] != ST_NONE)) {
    const uint8_t st = head[at(s, r)];
    if (st == ST_NONE) break;
    res.alignments.push_back(st);
    if (st == ST_INSERTION) ++s;
    else if (st == ST_DELETION) ++r;
    else {
      ++s;
      ++r;
    }
  }
  return res;
}

// ---------------------------------------------------------------------------------
// AssemblyUtils.discoverHaplotypes (M/assembly/AssemblyUtils.scala:50-114)
// ---------------------------------------------------------------------------------
inline std::vector<std::string> discoverHaplotypes(const std::vector<const MappedRead*>& reads, const std::string& currentReference, int kmerSize,
                                                   int minOccurrence, int minMeanKmerQuality, int expectedPloidy = 2,
                                                   int maxPathsToScore = 16) {
  auto paths = discoverPathSequences(reads, currentReference, kmerSize, minOccurrence, maxPathsToScore + 1, minMeanKmerQuality);
  if (paths.empty()) return {};
  if ((int)paths.size() <= expectedPloidy) return paths;
  if ((int)paths.size() <= maxPathsToScore) {
    std::vector<std::pair<long, std::string>> scored;
    for (auto& path : paths) {
      long total = 0;
      for (auto* read : reads) total += -(long)readAlignment(read->sequence, path).alignmentScore;  // :97-101
      scored.push_back({total, path});
    }
    std::stable_sort(scored.begin(), scored.end(), [](const std::pair<long, std::string>& a, const std::pair<long, std::string>& b) { return -a.first < -b.first; });
    std::vector<std::string> out;
    for (int i = 0; i < expectedPloidy && i < (int)scored.size(); ++i) out.push_back(scored[i].second);
    return out;
  }
  return {};  // too many paths: all variants skipped (:108-112)
}

struct AssemblyVariant {
  int64_t locus;
  std::string ref, alt;
  bool operator<(const AssemblyVariant& o) const { return std::tie(locus, ref, alt) < std::tie(o.locus, o.ref, o.alt); }
  bool operator==(const AssemblyVariant& o) const { return locus == o.locus && ref == o.ref && alt == o.alt; }
};

// AssemblyUtils.buildVariantsFromPath (:127-174)
inline std::vector<AssemblyVariant> buildVariantsFromPath(const std::string& path, int referenceStart, const ContigSequence& referenceContig,
                                                          const std::string& currentReference, bool allowReferenceVariant = false) {
  ReadAlignmentResult alignment = readAlignment(path, currentReference);
  // run-length encode into cigar elements
  struct El {
    uint8_t op;
    int len;
  };
  std::vector<El> cigar;
  for (uint8_t st : alignment.alignments) {
    if (!cigar.empty() && cigar.back().op == st) ++cigar.back().len;
    else cigar.push_back({st, 1});
  }
  int referenceIndex = alignment.refStartIdx, pathIndex = 0;
  std::vector<AssemblyVariant> out;
  for (size_t ci = 0; ci < cigar.size(); ++ci) {
    const uint8_t op = cigar[ci].op;
    const int referenceLength = (op == ST_INSERTION) ? 0 : cigar[ci].len;  // = X D consume reference
    const int pathLength = (op == ST_DELETION) ? 0 : cigar[ci].len;        // = X I consume the path
    const int64_t locus = (int64_t)referenceStart + referenceIndex;
    if (op == ST_MISMATCH) {
      std::string referenceAllele = referenceContig.slice(locus, locus + referenceLength);
      std::string alternateAllele = path.substr(pathIndex, pathLength);
      if (!referenceAllele.empty() && !alternateAllele.empty() && (allowReferenceVariant || referenceAllele != alternateAllele))
        out.push_back({locus, referenceAllele, alternateAllele});
    } else if ((op == ST_INSERTION || op == ST_DELETION) && ci != 0 && ci != cigar.size() - 1) {
      std::string referenceAllele = referenceContig.slice(locus - 1, locus + referenceLength);
      std::string alternateAllele = path.substr(std::max(0, pathIndex - 1), pathLength + 1);
      if (!referenceAllele.empty() && !alternateAllele.empty() && (allowReferenceVariant || referenceAllele != alternateAllele))
        out.push_back({locus - 1, referenceAllele, alternateAllele});
    }
    referenceIndex += referenceLength;
    pathIndex += pathLength;
  }
  return out;
}

}  // namespace guac_oracle
