// STAND-IN for the third-party header kseq++/seqio.hpp (https://github.com/cartoonist/kseqpp; an unpinned git submodule
// of the reference that is absent from this image), which the reference's include/reads.h includes.  A maintainer's
// build has the real library on its include path; this file only exists so that the reference's sources compile and
// run HERE.  It restates the one thing src/reads.cpp:3-18 uses: a FASTA/FASTQ stream (plain or gzip) whose read()
// returns every record with its `seq` — kseq's record grammar: '>' or '@' header line, sequence lines up to a line
// that starts with '+' (FASTQ) or with the next header, then as many quality characters as sequence characters.
#pragma once
#include <zlib.h>

#include <stdexcept>
#include <string>
#include <vector>
namespace klibpp {
struct KSeq {
  std::string name, comment, seq, qual;
};
class SeqStreamIn {
 public:
  explicit SeqStreamIn(const char *path) : f_(gzopen(path, "rb")) {
    if (!f_) throw std::runtime_error(std::string("cannot open ") + path);
  }
  ~SeqStreamIn() {
    if (f_) gzclose(f_);
  }
  SeqStreamIn(const SeqStreamIn &) = delete;
  SeqStreamIn &operator=(const SeqStreamIn &) = delete;
  std::vector<KSeq> read() {
    std::vector<KSeq> out;
    std::string line;
    bool have = next_line(line);
    while (have) {
      if (line.empty() || (line[0] != '>' && line[0] != '@')) {  // skip to the next header
        have = next_line(line);
        continue;
      }
      KSeq r;
      const size_t sp = line.find_first_of(" \t");
      r.name = line.substr(1, sp == std::string::npos ? std::string::npos : sp - 1);
      if (sp != std::string::npos) r.comment = line.substr(sp + 1);
      have = next_line(line);
      while (have && !line.empty() && line[0] != '+' && line[0] != '>' && line[0] != '@') {
        r.seq += line;
        have = next_line(line);
      }
      if (have && !line.empty() && line[0] == '+') {
        while (r.qual.size() < r.seq.size() && (have = next_line(line))) r.qual += line;
        have = next_line(line);
      }
      out.push_back(std::move(r));
    }
    return out;
  }

 private:
  bool next_line(std::string &line) {
    line.clear();
    char buf[1 << 16];
    for (;;) {
      if (!gzgets(f_, buf, sizeof buf)) return !line.empty();
      line += buf;
      if (!line.empty() && line.back() == '\n') {
        line.pop_back();
        if (!line.empty() && line.back() == '\r') line.pop_back();
        return true;
      }
    }
  }
  gzFile f_;
};
}  // namespace klibpp
