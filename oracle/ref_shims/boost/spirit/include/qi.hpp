// TEST INFRASTRUCTURE ONLY -- Boost.Spirit cannot be stubbed, so the reference's three Spirit GRAMMARS (FASTA alignment,
// loci list, value vector; none of them on the hot path) are replaced by hand-written recursive-descent parsers of the same
// class names and template signatures (the reference classes befriend those names), and qi::phrase_parse() forwards to them.
// The reference's own parse_Alignment / parse_Loci_list / parse_ValueVector (Alignment_parsers.hpp, Loci_parsers.hpp,
// ValueVector_parser.hpp) compile and run unmodified on top of this.
#pragma once
#define APEGRUNT_PARSER_COMMONS_HPP
#define APEGRUNT_ALIGNMENT_PARSER_FASTA_GRAMMAR_HPP
#define APEGRUNT_STATEVECTOR_PARSER_FASTA_GRAMMAR_HPP
#define APEGRUNT_LOCI_PARSER_GRAMMAR_HPP
#define APEGRUNT_VALUE_LIST_PARSER_GRAMMAR_HPP
#include <cctype>
#include <cstddef>
#include <cstdlib>
#include <istream>
#include <iterator>
#include <memory>
#include <string>
namespace boost { namespace spirit {
namespace ascii { struct space_type {}; static const space_type space{}; }
// single-pass stream iterator (only used for the .gz/.xz branch, which this build cannot decompress)
using istream_iterator = std::istreambuf_iterator<char>;
namespace qi {
template <class It, class Grammar, class Skipper, class Attr>
bool phrase_parse(It& first, It last, Grammar& grammar, const Skipper&, Attr& attr) { return grammar.parse(first, last, attr); }
}
namespace karma {}
}
namespace phoenix {}
}
namespace apegrunt {
namespace parsers {
inline void printer(const std::string&) {}
template <class It> inline void skip_space(It& p, It last) { while (p != last && std::isspace((unsigned char)*p)) ++p; }
template <class It> inline bool parse_uint(It& p, It last, std::size_t& v) {
    if (p == last || !std::isdigit((unsigned char)*p)) return false;
    v = 0; while (p != last && std::isdigit((unsigned char)*p)) { v = v * 10 + std::size_t(*p - '0'); ++p; } return true;
}

// '>' name-to-end-of-line, then the states of the record in chunks (whitespace skipped) -- Alignment_parser_FASTA_grammar.hpp:68-84,
// StateVector_parser_FASTA_grammar.hpp:60-68 (16k chunks handed to statevector_t::append)
template <typename IteratorT, typename StateVectorT> struct StateVector_parser_FASTA_grammar {          // the class the state vector befriends
    static void append(StateVectorT* seq, const std::string& chunk) { seq->append(chunk); }
};
template <typename IteratorT, typename AlignmentT> struct Alignment_parser_FASTA_grammar {
    using statevector_t = typename AlignmentT::statevector_t;
    template <class Attr> bool parse(IteratorT& p, IteratorT last, Attr& out) {
        auto alignment = std::make_shared<AlignmentT>();
        statevector_t* seq = nullptr;
        skip_space(p, last);
        while (p != last && *p == '>') {
            ++p;
            std::string name; while (p != last && *p != '\n') { name.push_back(*p); ++p; }
            if (p != last) ++p;
            seq = alignment->get_new_sequence(name);
            std::string chunk; chunk.reserve(16384);
            for (;;) {
                skip_space(p, last);
                if (p == last || *p == '>') break;
                chunk.push_back(*p); ++p;
                if (chunk.size() == 16384) { StateVector_parser_FASTA_grammar<IteratorT, statevector_t>::append(seq, chunk); chunk.clear(); }
            }
            if (!chunk.empty()) StateVector_parser_FASTA_grammar<IteratorT, statevector_t>::append(seq, chunk);
        }
        if (seq) alignment->set_nloci(seq->size());
        alignment->finalize();
        out = alignment;
        return true;
    }
};

// whitespace-separated "uint" or "uint-uint" entries -- Loci_parser_grammar.hpp:65-84
template <typename IteratorT, typename LociT> struct Loci_parser_grammar {
    typedef void (LociT::*push_back_one)(std::size_t locus, std::size_t base);                           // the two overloads are ambiguous
    typedef void (LociT::*push_back_range)(std::size_t locus_begin, std::size_t locus_end, std::size_t base);   // without the cast (as in the reference)
    explicit Loci_parser_grammar(std::size_t base_index = 1) : m_base_index(base_index) {}
    template <class Attr> bool parse(IteratorT& p, IteratorT last, Attr& out) {
        auto loci = std::make_shared<LociT>();
        for (;;) {
            skip_space(p, last);
            IteratorT save = p; std::size_t a = 0, b = 0;
            if (!parse_uint(p, last, a)) { p = save; break; }
            IteratorT after = p; skip_space(p, last);
            if (p != last && *p == '-') {
                ++p; skip_space(p, last);
                if (parse_uint(p, last, b)) { ((*loci).*static_cast<push_back_range>(&LociT::push_back))(a, b, m_base_index); continue; }
                p = save; break;                              // "12-" matches neither alternative
            }
            p = after; ((*loci).*static_cast<push_back_one>(&LociT::push_back))(a, m_base_index);
        }
        out = loci;
        return true;
    }
    std::size_t m_base_index;
};

// whitespace-separated doubles -- ValueVector_parser_grammar.hpp:64-76
template <typename IteratorT, typename ValueT> struct ValueVector_parser_grammar {
    template <class Attr> bool parse(IteratorT& p, IteratorT last, Attr& out) {
        auto values = std::make_shared<typename Attr::element_type>();     // ValueVector<ValueT> = std::vector<ValueT> (ValueVector_forward.h)
        for (;;) {
            skip_space(p, last); if (p == last) break;
            std::string tok; IteratorT q = p;
            while (q != last && !std::isspace((unsigned char)*q)) { tok.push_back(*q); ++q; }
            char* endp = nullptr; const double v = std::strtod(tok.c_str(), &endp);
            if (endp == tok.c_str()) break;
            values->push_back(ValueT(v)); p = q;
        }
        out = values;
        return true;
    }
};
}}
