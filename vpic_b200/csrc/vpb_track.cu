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
//  * the protected range is rounded OUT to page boundaries; a write to the few bytes of a neighbouring
//    allocation that share the first or last page also faults and merely costs one spurious upload;
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
struct Tracked { uintptr_t base; uintptr_t lo, hi; volatile int armed; volatile int dirty; };
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
      // a neighbouring array that shares a page with this one just lost the protection of that page:
      // it can no longer vouch for itself either
      for( int j=0; j<n; j++ ) {
        Tracked & o = g_tr[j];
        if( o.armed && o.lo<t.hi && t.lo<o.hi ) { o.dirty = 1; o.armed = 0; mprotect( (void *)o.lo, o.hi-o.lo, PROT_READ|PROT_WRITE ); }
      }
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
  const uintptr_t page = (uintptr_t)sysconf( _SC_PAGESIZE );
  Tracked * t = find( p );
  if( !t ) {
    if( g_ntr>=TR_MAX ) return;
    t = &g_tr[g_ntr]; t->base = (uintptr_t)p; t->armed = 0; t->dirty = 1; t->lo = t->hi = 0;
    __sync_synchronize();
    g_ntr = g_ntr+1;
  }
  const uintptr_t lo = (uintptr_t)p & ~( page-1 ), hi = ( (uintptr_t)p + bytes + page-1 ) & ~( page-1 );
  if( t->armed && ( t->lo!=lo || t->hi!=hi ) ) { mprotect( (void *)t->lo, t->hi-t->lo, PROT_READ|PROT_WRITE ); t->armed = 0; }
  t->lo = lo; t->hi = hi; t->dirty = 0;
  __sync_synchronize();
  if( !t->armed ) {
    t->armed = 1;
    if( mprotect( (void *)lo, hi-lo, PROT_READ )!=0 ) { t->armed = 0; t->dirty = 1; }   // not protectable: behave like STRICT
  }
}

// true: the host program has not written the array since it was armed (the device mirror is still equal to it)
bool track_clean( const void * p, size_t bytes ) {
  if( !g_on || !p ) return false;
  Tracked * t = find( p );
  if( !t || !t->armed || t->dirty ) return false;
  const uintptr_t page = (uintptr_t)sysconf( _SC_PAGESIZE );
  const uintptr_t hi = ( (uintptr_t)p + bytes + page-1 ) & ~( page-1 );
  return hi<=t->hi;                        // (an array that grew in place is not covered)
}

// The library itself is about to write the host array with CPU stores, or the array is going away.
void track_disarm( const void * p ) {
  Tracked * t = p ? find( p ) : nullptr;
  if( !t ) return;
  if( t->armed ) {
    t->armed = 0; mprotect( (void *)t->lo, t->hi-t->lo, PROT_READ|PROT_WRITE );
    for( int j=0; j<g_ntr; j++ ) {          // neighbours sharing a page (see on_segv)
      Tracked & o = g_tr[j];
      if( o.armed && o.lo<t->hi && t->lo<o.hi ) { o.dirty = 1; o.armed = 0; mprotect( (void *)o.lo, o.hi-o.lo, PROT_READ|PROT_WRITE ); }
    }
  }
  t->dirty = 1;
}

void track_forget_all() {
  for( int k=0; k<g_ntr; k++ ) if( g_tr[k].armed ) { g_tr[k].armed = 0; mprotect( (void *)g_tr[k].lo, g_tr[k].hi-g_tr[k].lo, PROT_READ|PROT_WRITE ); }
  g_ntr = 0;
}

} // namespace vpb
