// writers.cpp — Peng::printShortMeme / printJson (src/peng.cpp:602-727), byte for byte.
// Two properties of the reference's output are easy to miss and are kept: (1) `std::fixed << setprecision(8)` is set
// on the file stream while the FIRST matrix is written and never reset, so the header line of the first motif prints
// bg_prob / log(Pval) in the default format and every later motif prints them fixed with 8 decimals; (2)
// Utils::no_zero_pwm changes the matrix in place, so writing JSON after MEME applies it a second time.
#include <algorithm>
#include <fstream>
#include <iomanip>
#include <iostream>

#include "peng_host.h"

namespace peng_host {

static const unsigned kPrecision = 8;
static const char kAlphabet[] = "ACGT";

void write_meme(std::vector<Motif*>& motifs, const std::string& path, const float* v0) {
  std::sort(motifs.begin(), motifs.end(), by_log_pvalue);
  std::ofstream f(path);
  if (!f.is_open()) {
    std::cerr << "Unable to open output file (" << path << ")!";
    return;
  }
  f << "MEME version 4" << std::endl << std::endl;
  f << "ALPHABET= " << kAlphabet << std::endl << std::endl;
  f << "Background letter frequencies" << std::endl;
  for (int i = 0; i < 4; i++) {
    if (i) f << " ";
    f << kAlphabet[i] << " " << v0[i];
  }
  f << std::endl << std::endl;
  for (Motif* m : motifs) {
    f << "MOTIF " << pwm_string(m->pwm) << std::endl;
    f << "letter-probability matrix:"
      << " alength= " << 4 << " w= " << (size_t)m->length << " nsites= " << m->n_sites << " bg_prob= " << m->bg_p
      << " opt_bg_order= " << m->opt_bg_order << " log(Pval)= " << m->log_pvalue << std::endl;
    no_zero_pwm(m->pwm, kPrecision);
    for (const PwmRow& row : m->pwm) {
      for (int a = 0; a < 4; a++) {
        if (a) f << " ";
        f << std::fixed << std::setprecision(kPrecision) << row[a];
      }
      f << std::endl;
    }
    f << std::endl;
  }
}

void write_json(std::vector<Motif*>& motifs, const std::string& path, const float* v0) {
  std::sort(motifs.begin(), motifs.end(), by_log_pvalue);
  std::ofstream f(path);
  if (!f.is_open()) {
    std::cerr << "Unable to open output file (" << path << ")!";
    return;
  }
  f << "{" << std::endl;
  f << "\t\"alphabet\" : \"" << kAlphabet << "\"," << std::endl;
  f << "\t\"bg\" : [";
  for (int i = 0; i < 4; i++) {
    f << v0[i];
    if (i != 3) f << ", ";
  }
  f << "]," << std::endl;
  f << "\t\"alphabet_length\" : " << 4 << "," << std::endl;
  f << "\t\"patterns\" : [" << std::endl;
  for (Motif* m : motifs) {
    f << "\t\t{" << std::endl;
    f << "\t\t\t\"iupac_motif\" : \"" << pwm_string(m->pwm) << "\"," << std::endl;
    f << "\t\t\t\"pattern_length\" : " << (size_t)m->length << "," << std::endl;
    f << "\t\t\t\"sites\" : " << m->n_sites << "," << std::endl;
    f << "\t\t\t\"log(Pval)\" : " << m->log_pvalue << "," << std::endl;
    f << "\t\t\t\"bg_prob\" : " << m->bg_p << "," << std::endl;
    f << "\t\t\t\"opt_bg_order\" : " << m->opt_bg_order << "," << std::endl;
    f << "\t\t\t\"pwm\" : [" << std::endl;
    no_zero_pwm(m->pwm, kPrecision);
    for (size_t w = 0; w < m->pwm.size(); w++) {
      f << "\t\t\t\t\t[";
      for (int a = 0; a < 4; a++) {
        f << std::fixed << std::setprecision(kPrecision) << m->pwm[w][a];
        f << (a != 3 ? ", " : "]");
      }
      if (w + 1 != m->pwm.size()) f << ", ";
      f << std::endl;
    }
    f << "\t\t\t\t]" << std::endl;
    f << "\t\t}";
    if (m != motifs.back()) f << ",";
    f << std::endl;
  }
  f << "\t]" << std::endl;
  f << "}" << std::endl;
}

}  // namespace peng_host
