// gr_standin.h -- TEST-ONLY stand-in for the slice of GNU Radio 3.10 the shims touch (gr::block, gr::tagged_stream_block,
// gr::tag_t, pmt).  GNU Radio is not installed in this image; when find_package(Gnuradio) succeeds (CMakeLists.txt) the shims
// are compiled against the real headers and this file is not used.  It is NOT a scheduler: buffers are whatever the test
// driver passes to work(), tags live in a per-port vector, messages are appended to a per-port log.
#pragma once
#include <complex>
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

typedef std::complex<float> gr_complex;
typedef std::vector<int> gr_vector_int;
typedef std::vector<const void*> gr_vector_const_void_star;
typedef std::vector<void*> gr_vector_void_star;

namespace pmt {
struct pmt_base {
    enum kind_t { SYMBOL, LONG, DOUBLE, C32VEC, DICT, NIL } kind = NIL;
    std::string sym; long l = 0; double d = 0; std::vector<gr_complex> v;
    std::vector<std::pair<std::shared_ptr<pmt_base>, std::shared_ptr<pmt_base>>> items;
};
typedef std::shared_ptr<pmt_base> pmt_t;
inline pmt_t string_to_symbol(const std::string& s)                 // interned, so that key == key compares as in pmt
{
    static std::map<std::string, pmt_t> table;
    pmt_t& p = table[s];
    if (!p) { p = std::make_shared<pmt_base>(); p->kind = pmt_base::SYMBOL; p->sym = s; }
    return p;
}
inline pmt_t mp(const std::string& s) { return string_to_symbol(s); }
inline std::string symbol_to_string(const pmt_t& p) { if (!p || p->kind != pmt_base::SYMBOL) throw std::runtime_error("pmt: not a symbol"); return p->sym; }
inline pmt_t from_long(long x) { auto p = std::make_shared<pmt_base>(); p->kind = pmt_base::LONG; p->l = x; return p; }
inline pmt_t from_double(double x) { auto p = std::make_shared<pmt_base>(); p->kind = pmt_base::DOUBLE; p->d = x; return p; }
inline long to_long(const pmt_t& p) { if (!p || p->kind != pmt_base::LONG) throw std::runtime_error("pmt: not a long"); return p->l; }
inline double to_double(const pmt_t& p)
{
    if (p && p->kind == pmt_base::LONG) return (double)p->l;
    if (!p || p->kind != pmt_base::DOUBLE) throw std::runtime_error("pmt: not a double");
    return p->d;
}
inline pmt_t init_c32vector(size_t n, const std::vector<gr_complex>& d)
{
    auto p = std::make_shared<pmt_base>(); p->kind = pmt_base::C32VEC; p->v.assign(d.begin(), d.begin() + n); return p;
}
inline std::vector<gr_complex> c32vector_elements(const pmt_t& p) { if (!p || p->kind != pmt_base::C32VEC) throw std::runtime_error("pmt: not a c32vector"); return p->v; }
inline pmt_t make_dict() { auto p = std::make_shared<pmt_base>(); p->kind = pmt_base::DICT; return p; }
inline pmt_t dict_add(const pmt_t& d, const pmt_t& k, const pmt_t& v)
{
    auto p = std::make_shared<pmt_base>(*d);
    for (auto& it : p->items) if (it.first == k) { it.second = v; return p; }
    p->items.push_back({ k, v });
    return p;
}
inline pmt_t dict_ref(const pmt_t& d, const pmt_t& k, const pmt_t& dflt) { for (auto& it : d->items) if (it.first == k) return it.second; return dflt; }
} // namespace pmt

namespace gr {
struct tag_t { uint64_t offset = 0; pmt::pmt_t key, value, srcid; };

struct io_signature {
    int n, itemsize;
    static std::shared_ptr<io_signature> make(int mn, int mx, int sz) { (void)mx; return std::make_shared<io_signature>(io_signature{ mn, sz }); }
    static std::shared_ptr<io_signature> make2(int mn, int mx, int s1, int s2) { (void)mx; (void)s2; return std::make_shared<io_signature>(io_signature{ mn, s1 }); }
};

class block
{
public:
    enum tag_propagation_policy_t { TPP_DONT = 0, TPP_ALL_TO_ALL = 1, TPP_ONE_TO_ONE = 2 };
    block(const std::string& name, std::shared_ptr<io_signature> in, std::shared_ptr<io_signature> out)
        : in_tags(in->n), out_tags(out->n), d_name(name), d_read(in->n, 0), d_written(out->n, 0) {}
    virtual ~block() {}
    virtual void forecast(int, gr_vector_int& req) { for (auto& r : req) r = 1; }
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
                             gr_vector_void_star& output_items) = 0;
    const std::string& name() const { return d_name; }
    uint64_t nitems_read(unsigned i) const { return d_read[i]; }
    uint64_t nitems_written(unsigned i) const { return d_written[i]; }
    void get_tags_in_range(std::vector<tag_t>& v, unsigned port, uint64_t a, uint64_t b) const
    {
        v.clear();
        for (auto& t : in_tags[port]) if (t.offset >= a && t.offset < b) v.push_back(t);
    }
    void get_tags_in_window(std::vector<tag_t>& v, unsigned port, uint64_t a, uint64_t b) const { get_tags_in_range(v, port, d_read[port] + a, d_read[port] + b); }
    void add_item_tag(unsigned port, uint64_t off, const pmt::pmt_t& k, const pmt::pmt_t& val) { tag_t t; t.offset = off; t.key = k; t.value = val; out_tags[port].push_back(t); }
    void consume_each(int n) { d_consumed = n; }
    void message_port_register_out(const pmt::pmt_t&) {}
    void message_port_pub(const pmt::pmt_t& port, const pmt::pmt_t& msg) { messages[pmt::symbol_to_string(port)].push_back(msg); }
    void set_tag_propagation_policy(tag_propagation_policy_t) {}
    void set_max_noutput_items(int) {}
    void set_min_noutput_items(int) {}
    void set_output_multiple(int) {}
    void set_relative_rate(uint64_t, uint64_t) {}
    // ---- what the scheduler does around a work() call, for the test driver
    std::vector<std::vector<tag_t>> in_tags, out_tags;
    std::map<std::string, std::vector<pmt::pmt_t>> messages;
    int run(int noutput_items, gr_vector_int& nin, gr_vector_const_void_star& in, gr_vector_void_star& out)
    {
        d_consumed = -1;
        const int produced = general_work(noutput_items, nin, in, out);
        const int consumed = d_consumed >= 0 ? d_consumed : (nin.empty() ? 0 : nin[0]);      // sync/tagged-stream blocks consume what they were given
        for (auto& r : d_read) r += consumed;
        if (produced > 0) for (auto& w : d_written) w += produced;
        for (auto& tv : in_tags) { std::vector<tag_t> keep; for (auto& t : tv) if (t.offset >= d_read[0]) keep.push_back(t); tv.swap(keep); }
        return produced;
    }
    int consumed() const { return d_consumed; }
protected:
    block() {}                                          // for the interface classes that inherit virtually (as gr::block has)
    std::string d_name;
    std::vector<uint64_t> d_read, d_written;
    int d_consumed = -1;
};

// one tagged stream (= one frame) per work() call: ninput_items[0] = the length tag's value
class tagged_stream_block : public block
{
public:
    tagged_stream_block(const std::string& name, std::shared_ptr<io_signature> in, std::shared_ptr<io_signature> out, const std::string& len_key)
        : block(name, in, out), d_length_tag_key_str(len_key) {}
    virtual int work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    int general_work(int n, gr_vector_int& nin, gr_vector_const_void_star& in, gr_vector_void_star& out) override
    {
        const int r = work(n, nin, in, out);
        if (r > 0) for (unsigned p = 0; p < out_tags.size(); p++) add_item_tag(p, nitems_written(p), pmt::mp(d_length_tag_key_str), pmt::from_long(r));   // update_length_tags
        return r;
    }
protected:
    tagged_stream_block() {}
    std::string d_length_tag_key_str;
};
} // namespace gr

namespace gnuradio {
template <class T, class... A> std::shared_ptr<T> make_block_sptr(A&&... a) { return std::make_shared<T>(std::forward<A>(a)...); }
}
