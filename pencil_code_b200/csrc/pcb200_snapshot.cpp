// pcb200_snapshot.cpp -- var.dat snapshots of one rank in the reference's distributed-I/O layout.
//
// The reference writes data/procN/var.dat with two Fortran sequential unformatted records
// (src/io_dist.f90:108-199 output_snap, io_strategy "dist"):
//   record 1: a(1:mx,1:my,1:mz,1:nvar)                      (binary64, x fastest: the host f-array as it is)
//   record 2: t, x(1:mx), y(1:my), z(1:mz), dx, dy, dz      (binary64 with REAL_PRECISION=double; no shear)
// and reads them back in read_snap (src/io_dist.f90:697-788, called from input_snap :657-695).
// Record framing is gfortran's (the compiler of every reference config): a 4-byte length before and after the
// payload; a record longer than 2147483639 bytes is cut into subrecords, the leading marker of a subrecord is
// negative when another subrecord follows, the trailing marker is negative when one precedes.  A 512^3 snapshot
// (8.9 GB) therefore has 5 subrecords.  This file is host-only code: it touches neither the device nor the oracle.
#include "../../include/pcb200.h"

#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <fcntl.h>
#include <unistd.h>

namespace {

long long subrecord_max() {
  const char *e = std::getenv("PCB200_SUBRECORD_MAX"); // test hook: exercise the subrecord framing with small files
  if (e && std::atoll(e) > 0) return std::atoll(e);
  return 2147483639LL;
}

bool write_all(int fd, const void *p, size_t n) {
  const char *c = static_cast<const char *>(p);
  while (n) {
    const size_t chunk = n < (size_t(1) << 30) ? n : (size_t(1) << 30);
    const ssize_t w = ::write(fd, c, chunk);
    if (w < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    c += w;
    n -= (size_t)w;
  }
  return true;
}

bool read_all(int fd, void *p, size_t n) {
  char *c = static_cast<char *>(p);
  while (n) {
    const size_t chunk = n < (size_t(1) << 30) ? n : (size_t(1) << 30);
    const ssize_t r = ::read(fd, c, chunk);
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    if (r == 0) return false; // short file
    c += r;
    n -= (size_t)r;
  }
  return true;
}

// one Fortran record made of the given pieces (written back to back as one payload)
struct Piece {
  const void *p;
  size_t n;
};

bool write_record(int fd, const std::vector<Piece> &pieces) {
  size_t total = 0;
  for (const Piece &q : pieces) total += q.n;
  const size_t smax = (size_t)subrecord_max();
  size_t done = 0, ip = 0, off = 0; // position: piece ip, byte off
  bool first = true;
  do {
    const size_t len = total - done < smax ? total - done : smax;
    const bool more = done + len < total;
    const int32_t lead = more ? -(int32_t)len : (int32_t)len;
    const int32_t trail = first ? (int32_t)len : -(int32_t)len;
    if (!write_all(fd, &lead, 4)) return false;
    size_t left = len;
    while (left) {
      const size_t n = pieces[ip].n - off < left ? pieces[ip].n - off : left;
      if (!write_all(fd, static_cast<const char *>(pieces[ip].p) + off, n)) return false;
      off += n;
      left -= n;
      if (off == pieces[ip].n) { ++ip; off = 0; }
    }
    if (!write_all(fd, &trail, 4)) return false;
    done += len;
    first = false;
  } while (done < total);
  return true;
}

struct MutPiece {
  void *p;
  size_t n;
};

// returns 0, or 2 for a framing/length mismatch, 1 for an I/O error
int read_record(int fd, const std::vector<MutPiece> &pieces) {
  size_t total = 0;
  for (const MutPiece &q : pieces) total += q.n;
  size_t done = 0, ip = 0, off = 0;
  bool first = true;
  for (;;) {
    int32_t lead;
    if (!read_all(fd, &lead, 4)) return 1;
    const bool more = lead < 0;
    const size_t len = (size_t)(more ? -(int64_t)lead : (int64_t)lead);
    if (done + len > total) return 2;
    size_t left = len;
    while (left) {
      const size_t n = pieces[ip].n - off < left ? pieces[ip].n - off : left;
      if (!read_all(fd, static_cast<char *>(pieces[ip].p) + off, n)) return 1;
      off += n;
      left -= n;
      if (off == pieces[ip].n) { ++ip; off = 0; }
    }
    int32_t trail;
    if (!read_all(fd, &trail, 4)) return 1;
    if ((size_t)(trail < 0 ? -(int64_t)trail : (int64_t)trail) != len || (trail < 0) == first) return 2;
    done += len;
    first = false;
    if (!more) break;
  }
  return done == total ? 0 : 2;
}

} // namespace

extern "C" int pcb200_write_snapshot(const char *path, const double *f, int mx, int my, int mz, int nvar, double t,
                                     const double *x, const double *y, const double *z, double dx, double dy, double dz) {
  if (!path || !f || !x || !y || !z || mx <= 0 || my <= 0 || mz <= 0 || nvar <= 0) return 3;
  const int fd = ::open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644); // output_snap deletes an existing file first
  if (fd < 0) return 1;
  const size_t nf = (size_t)mx * my * mz * nvar * sizeof(double);
  bool ok = write_record(fd, {{f, nf}});
  const double d3[3] = {dx, dy, dz};
  ok = ok && write_record(fd, {{&t, 8}, {x, (size_t)mx * 8}, {y, (size_t)my * 8}, {z, (size_t)mz * 8}, {d3, 24}});
  if (::close(fd) != 0) ok = false;
  return ok ? 0 : 1;
}

extern "C" int pcb200_read_snapshot(const char *path, double *f, int mx, int my, int mz, int nvar, double *t, double *x,
                                    double *y, double *z, double *dxyz) {
  if (!path || !f || !t || !x || !y || !z || !dxyz || mx <= 0 || my <= 0 || mz <= 0 || nvar <= 0) return 3;
  const int fd = ::open(path, O_RDONLY);
  if (fd < 0) return 1;
  const size_t nf = (size_t)mx * my * mz * nvar * sizeof(double);
  int rc = read_record(fd, {{f, nf}});
  if (rc == 0) rc = read_record(fd, {{t, 8}, {x, (size_t)mx * 8}, {y, (size_t)my * 8}, {z, (size_t)mz * 8}, {dxyz, 24}});
  ::close(fd);
  return rc;
}
