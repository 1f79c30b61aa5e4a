// Minimal stand-in for the three Boost headers the reference smcsmc includes (test/oracle tooling only).
// zlib-backed so that <prefix>.recomb.gz is a real gzip file.  Not part of the product.
#ifndef SMC_BOOST_SHIM_FILTERING_STREAM
#define SMC_BOOST_SHIM_FILTERING_STREAM
#include <zlib.h>
#include <istream>
#include <ostream>
#include <sstream>
#include <string>
#include <streambuf>
#include <vector>
namespace boost { namespace iostreams {
struct gzip_compressor {};
struct gzip_decompressor {};

class gz_obuf : public std::stringbuf {
 public:
  std::ostream* sink = nullptr;
  bool gz = false;
  ~gz_obuf() { finish(); }
  void finish() {
    if (!sink) return;
    std::string s = str();
    if (gz) {
      z_stream zs; zs.zalloc = Z_NULL; zs.zfree = Z_NULL; zs.opaque = Z_NULL;
      deflateInit2(&zs, Z_DEFAULT_COMPRESSION, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY);
      zs.next_in = (Bytef*)s.data(); zs.avail_in = (uInt)s.size();
      std::vector<unsigned char> buf(1 << 16);
      int ret;
      do {
        zs.next_out = buf.data(); zs.avail_out = (uInt)buf.size();
        ret = deflate(&zs, Z_FINISH);
        sink->write((const char*)buf.data(), buf.size() - zs.avail_out);
      } while (ret != Z_STREAM_END);
      deflateEnd(&zs);
    } else {
      sink->write(s.data(), s.size());
    }
    sink->flush();
    sink = nullptr;
  }
};

class filtering_ostream : public std::ostream {
  gz_obuf buf_;
 public:
  filtering_ostream() : std::ostream(&buf_) {}
  ~filtering_ostream() { buf_.finish(); }
  void push(const gzip_compressor&) { buf_.gz = true; }
  void push(std::ostream& os) { buf_.sink = &os; }
};

class filtering_istream : public std::istream {
  std::stringbuf buf_;
  bool gz_ = false;
 public:
  filtering_istream() : std::istream(&buf_) {}
  void push(const gzip_decompressor&) { gz_ = true; }
  void push(std::istream& is) {
    std::stringstream raw; raw << is.rdbuf();
    std::string s = raw.str();
    if (!gz_) { buf_.str(s); return; }
    z_stream zs; zs.zalloc = Z_NULL; zs.zfree = Z_NULL; zs.opaque = Z_NULL;
    inflateInit2(&zs, 15 + 32);
    zs.next_in = (Bytef*)s.data(); zs.avail_in = (uInt)s.size();
    std::string out; std::vector<unsigned char> b(1 << 16);
    int ret;
    do {
      zs.next_out = b.data(); zs.avail_out = (uInt)b.size();
      ret = inflate(&zs, Z_NO_FLUSH);
      out.append((const char*)b.data(), b.size() - zs.avail_out);
    } while (ret == Z_OK);
    inflateEnd(&zs);
    buf_.str(out);
  }
};
}}
#endif
