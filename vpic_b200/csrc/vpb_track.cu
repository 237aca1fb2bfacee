// vpb_track.cu -- write tracking of the host program's arrays (coherence mode TRACKED).
//
// STRICT coherence has to assume that the host program wrote its arrays between any two calls, so
// every entry point uploads what it reads -- 12.4 GB per step on the e2e benchmark, of which only the
// first upload of each array ever carried new data.  TRACKED is STRICT plus knowledge: after an upload
// or a download the host array and the device mirror are identical, and the array's pages are made
// read-only (mprotect PROT_READ).  A later host WRITE faults; the SIGSEGV handler gives the pages back
// (PROT_READ|PROT_WRITE), marks the array dirty and returns, so the store is simply retried.  The next
// entry point uploads the dirty arrays and skips the clean ones.  Reads are never intercepted: the
// host copy is kept current exactly as in STRICT mode (downloads stay), only the redundant uploads go.
//
// Scope and limits (why it is a mode a host opts into, not the default):
//  * only the pages that lie ENTIRELY inside the array are protected.  The head and the tail of an array that is
//    not page aligned share their pages with whatever the allocator put next to it (another array, malloc's own
//    bookkeeping), so those two fragments (< 4 KB each) stay writable and are compared against a copy taken when
//    the array was armed instead.  A neighbour's stores therefore never fault and never mark this array dirty;
//  * stores are intercepted, system calls are not: a read()/fread() INTO a protected array fails with
//    EFAULT instead of faulting.  A host program that fills its arrays that way (restart files are
//    read into fresh arrays, which are not yet tracked) must stay STRICT or call vpic_b200_forget first;
//  * arrays below 1 MB are not tracked (they are always uploaded, as in STRICT mode);
//  * DMA writes of our own downloads are not affected by page protection; small copies that the CUDA
//    driver stages through the CPU fault like any other store and only cost a spurious upload.
#include "vpb_runtime.h"
#include <signal.h>
#include <sys/mman.h>
#include <unistd.h>
#include <cstring>

namespace vpb {

namespace {

constexpr int TR_MAX = 256;
struct Tracked {
  uintptr_t base; size_t bytes;       // the array as it was armed
  uintptr_t lo, hi;                   // its interior pages (protected while armed)
  volatile int armed; volatile int dirty;
  unsigned char * edge;               // copies of [base,lo) and [hi,base+bytes)
};
Tracked g_tr[TR_MAX];
volatile int g_ntr = 0;
bool g_on = false, g_handler = false;
struct sigaction g_old;

void on_segv( int sig, siginfo_t * si, void * ctx ) {
  const uintptr_t a = (uintptr_t)si->si_addr;
  const int n = g_ntr;
  for( int k=0; k<n; k++ ) {
    Tracked & t = g_tr[k];
    if( t.armed && a>=t.lo && a<t.hi ) {
      t.dirty = 1; t.armed = 0;
      mprotect( (void *)t.lo, t.hi-t.lo, PROT_READ|PROT_WRITE );     // async-signal-safe
      return;                                                        // the faulting store is retried
    }
  }
  // not ours: hand over to whoever was there before, or die the default way
  if( g_old.sa_flags & SA_SIGINFO ) { if( g_old.sa_sigaction ) { g_old.sa_sigaction( sig, si, ctx ); return; } }
  else if( g_old.sa_handler!=SIG_DFL && g_old.sa_handler!=SIG_IGN ) { g_old.sa_handler( sig ); return; }
  signal( SIGSEGV, SIG_DFL );
}

Tracked * find( const void * p ) {
  const uintptr_t b = (uintptr_t)p;
  for( int k=0; k<g_ntr; k++ ) if( g_tr[k].base==b ) return &g_tr[k];
  return nullptr;
}

uintptr_t page_size() { static const uintptr_t page = (uintptr_t)sysconf( _SC_PAGESIZE ); return page; }

void unprotect( Tracked * t ) {
  if( t->armed ) { t->armed = 0; mprotect( (void *)t->lo, t->hi-t->lo, PROT_READ|PROT_WRITE ); }
}

} // namespace

bool track_enabled() { return g_on; }

void track_enable( bool on ) {
  if( on && !g_handler ) {
    struct sigaction sa; memset( &sa, 0, sizeof(sa) );
    sa.sa_sigaction = on_segv; sa.sa_flags = SA_SIGINFO | SA_NODEFER;
    sigemptyset( &sa.sa_mask );
    if( sigaction( SIGSEGV, &sa, &g_old )!=0 ) VPB_ERROR( "cannot install the write-tracking fault handler" );
    g_handler = true;
  }
  if( !on ) track_forget_all();
  g_on = on;
}

// The host array at p (bytes long) and its device mirror are identical right now: arm the tracker.
void track_arm( const void * p, size_t bytes ) {
  if( !g_on || !p || bytes<( 1u<<20 ) ) return;
  const uintptr_t page = page_size();
  Tracked * t = find( p );
  if( !t ) {
    if( g_ntr>=TR_MAX ) return;
    t = &g_tr[g_ntr]; t->base = (uintptr_t)p; t->bytes = 0; t->armed = 0; t->dirty = 1; t->lo = t->hi = 0;
    t->edge = (unsigned char *)malloc( 2*page );
    if( !t->edge ) return;
    __sync_synchronize();
    g_ntr = g_ntr+1;
  }
  const uintptr_t b = (uintptr_t)p, lo = ( b + page-1 ) & ~( page-1 ), hi = ( b + bytes ) & ~( page-1 );
  if( t->armed && ( t->lo!=lo || t->hi!=hi ) ) unprotect( t );
  t->bytes = bytes; t->lo = lo; t->hi = hi; t->dirty = 0;
  memcpy( t->edge, (const void *)b, lo-b );
  memcpy( t->edge + page, (const void *)hi, b+bytes-hi );
  __sync_synchronize();
  if( !t->armed ) {
    t->armed = 1;
    if( mprotect( (void *)lo, hi-lo, PROT_READ )!=0 ) { t->armed = 0; t->dirty = 1; }   // not protectable: behave like STRICT
  }
}

// true: the host program has not written the first `bytes` of the array since it was armed (the device mirror is
// still equal to them)
bool track_clean( const void * p, size_t bytes ) {
  if( !g_on || !p ) return false;
  Tracked * t = find( p );
  if( !t || !t->armed || t->dirty || bytes>t->bytes ) return false;     // (an array that grew in place is not covered)
  const uintptr_t b = t->base, end = b + bytes;
  const size_t head = ( end<t->lo ? end : t->lo ) - b;
  if( memcmp( t->edge, (const void *)b, head )!=0 ) return false;
  if( end>t->hi && memcmp( t->edge + page_size(), (const void *)t->hi, end-t->hi )!=0 ) return false;
  return true;
}

// The library itself is about to write the host array with CPU stores, or the array is going away.
void track_disarm( const void * p ) {
  Tracked * t = p ? find( p ) : nullptr;
  if( !t ) return;
  unprotect( t );
  t->dirty = 1;
}

void track_forget_all() {
  for( int k=0; k<g_ntr; k++ ) { unprotect( &g_tr[k] ); free( g_tr[k].edge ); g_tr[k].edge = nullptr; }
  g_ntr = 0;
}

} // namespace vpb
