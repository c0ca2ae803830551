//! plonky2 (patched) `src/fri/oracle.rs` -- `PolynomialBatch` backed by libnlc_commit.so.
//!
//! Same public fields and signatures as upstream plonky2 0.1.4 @ d2598bd (pinned by the reference's Cargo.lock), so the
//! unchanged call sites keep compiling:
//!   vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93   PolynomialBatch::from_values(cols, rate_bits, false, cap_height, timing, None)
//!   vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:171-178 PolynomialBatch::from_coeffs(chunks, rate_bits, false, cap_height, timing, None)
//!   vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:200-206 PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params, timing)
//!   vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:289-294 commitment.get_lde_values_packed::<P>(i_start, step)
//!   vendor/starkyx/starkyx/src/plonky2/stark/proof.rs:307-310   commitment.polynomials (evaluated at zeta)
//! The workspace swaps it in the way it already swaps plonky2 (`/root/reference/Cargo.toml:64-70`, `[patch]`).
//!
//! NOT COMPILED IN THE GRAFT IMAGE (no rustc): this file is the transliteration of `include/nlc_plonky2.hpp`, whose
//! every call is exercised by `tests/cpp/test_plonky2_mirror.cpp` and, through ctypes, by the pytest suite.
//! `F = GoldilocksField` is `#[repr(transparent)]` over `u64`, so `&[F]` crosses the boundary as `*const u64`.

use std::sync::Arc;

use nlc_commit_sys as sys;

use crate::field::extension::Extendable;
use crate::field::packed::PackedField;
use crate::field::polynomial::{PolynomialCoeffs, PolynomialValues};
use crate::fri::proof::{FriInitialTreeProof, FriProof, FriQueryRound, FriQueryStep};
use crate::fri::structure::FriInstanceInfo;
use crate::fri::FriParams;
use crate::hash::hash_types::RichField;
use crate::hash::merkle_proofs::MerkleProof;
use crate::hash::merkle_tree::{MerkleCap, MerkleTree};
use crate::iop::challenger::Challenger;
use crate::plonk::config::{GenericConfig, Hasher};
use crate::util::fft::FftRootTable;
use crate::util::timing::TimingTree;
use crate::util::log2_strict;

/// One context per (thread, device): rayon workers and tokio blocking threads commit concurrently; calls on one
/// context serialise on its stream.
pub(crate) fn ctx() -> *mut sys::nlc_ctx {
    thread_local! {
        static CTX: *mut sys::nlc_ctx = {
            let mut c = std::ptr::null_mut();
            // device < 0: NLC_DEVICE, else the current CUDA device
            let rc = unsafe { sys::nlc_init(-1, &mut c) };
            assert!(rc == 0, "nlc_init failed: {}", last_error(std::ptr::null_mut()));
            c
        };
    }
    CTX.with(|c| *c)
}

fn last_error(c: *mut sys::nlc_ctx) -> String {
    unsafe { std::ffi::CStr::from_ptr(sys::nlc_last_error(c)).to_string_lossy().into_owned() }
}

/// plonky2's contract at this boundary is infallible-or-panic.
#[inline]
pub(crate) fn check(rc: i32) {
    if rc != 0 {
        panic!("nlc_commit: {}", last_error(ctx()));
    }
}

/// The device-resident batch; dropped with the last clone of the `PolynomialBatch` that owns it.
pub struct DeviceBatch(pub(crate) *mut sys::nlc_batch);
unsafe impl Send for DeviceBatch {}
unsafe impl Sync for DeviceBatch {}
impl Drop for DeviceBatch {
    fn drop(&mut self) {
        unsafe { sys::nlc_batch_free(self.0) }
    }
}

/// Represents a FRI oracle, i.e. a batch of polynomials which have been Merklized.
pub struct PolynomialBatch<F: RichField + Extendable<D>, C: GenericConfig<D, F = F>, const D: usize> {
    pub polynomials: Vec<PolynomialCoeffs<F>>,
    pub merkle_tree: MerkleTree<F, C::Hasher>,
    pub degree_log: usize,
    pub rate_bits: usize,
    pub blinding: bool,
    /// not in upstream: the HBM image (coefficients, column-major LDE in leaf order, digests, cap)
    pub(crate) handle: Arc<DeviceBatch>,
}

impl<F: RichField + Extendable<D>, C: GenericConfig<D, F = F>, const D: usize> PolynomialBatch<F, C, D> {
    /// Creates a list polynomial commitment for the polynomials interpolating the values in `values`.
    pub fn from_values(
        values: Vec<PolynomialValues<F>>,
        rate_bits: usize,
        blinding: bool,
        cap_height: usize,
        _timing: &mut TimingTree, // the scopes are NVTX ranges / nlc_timing_read on the device side
        _fft_root_table: Option<&FftRootTable<F>>,
    ) -> Self {
        assert!(!blinding, "blinding is false at every call site on this path (prover.rs:91-93)");
        assert!(!values.is_empty(), "PolynomialBatch: at least one polynomial required");
        let n = values[0].len();
        assert!(values.iter().all(|v| v.len() == n), "All polynomials in a batch must have the same length");
        let (w, log_n, big_n) = (values.len(), log2_strict(n), n << rate_bits);
        // The PUBLIC fields are filled while the commit runs (nlc_commit_from_values_materialised): upstream's own
        // serializer reads them, so they cannot be hidden behind accessors.
        let mut polynomials: Vec<PolynomialCoeffs<F>> = (0..w).map(|_| PolynomialCoeffs::new(vec![F::ZERO; n])).collect();
        let mut leaves_flat = vec![F::ZERO; if cfg!(feature = "lazy-merkle-fields") { 0 } else { big_n * w }];
        let num_digests = 2 * (big_n - (1 << cap_height));
        let mut digests = vec![<C::Hasher as Hasher<F>>::Hash::default(); if cfg!(feature = "lazy-merkle-fields") { 0 } else { num_digests }];
        let mut cap = vec![<C::Hasher as Hasher<F>>::Hash::default(); 1 << cap_height];
        let cols: Vec<*const u64> = values.iter().map(|v| v.values.as_ptr() as *const u64).collect();
        let coeff_out: Vec<*mut u64> = polynomials.iter_mut().map(|p| p.coeffs.as_mut_ptr() as *mut u64).collect();
        let mut h = std::ptr::null_mut();
        check(unsafe {
            sys::nlc_commit_from_values_materialised(
                ctx(), cols.as_ptr(), w, log_n as i32, rate_bits as i32, cap_height as i32, coeff_out.as_ptr(),
                if leaves_flat.is_empty() { std::ptr::null_mut() } else { leaves_flat.as_mut_ptr() as *mut u64 },
                if digests.is_empty() { std::ptr::null_mut() } else { digests.as_mut_ptr() as *mut u64 },
                cap.as_mut_ptr() as *mut u64, &mut h,
            )
        });
        Self {
            polynomials,
            merkle_tree: MerkleTree::from_flat_parts(leaves_flat, w, digests, MerkleCap(cap)),
            degree_log: log_n,
            rate_bits,
            blinding,
            handle: Arc::new(DeviceBatch(h)),
        }
    }

    /// Creates a list polynomial commitment for the polynomials `polynomials`.
    pub fn from_coeffs(
        polynomials: Vec<PolynomialCoeffs<F>>,
        rate_bits: usize,
        blinding: bool,
        cap_height: usize,
        _timing: &mut TimingTree,
        _fft_root_table: Option<&FftRootTable<F>>,
    ) -> Self {
        assert!(!blinding, "blinding is false at every call site on this path (prover.rs:171-178)");
        assert!(!polynomials.is_empty(), "PolynomialBatch: at least one polynomial required");
        let n = polynomials[0].len();
        assert!(polynomials.iter().all(|p| p.len() == n), "All polynomials in a batch must have the same length");
        let log_n = log2_strict(n);
        let cols: Vec<*const u64> = polynomials.iter().map(|p| p.coeffs.as_ptr() as *const u64).collect();
        let mut h = std::ptr::null_mut();
        check(unsafe {
            sys::nlc_commit_from_coeffs(ctx(), cols.as_ptr(), cols.len(), log_n as i32, rate_bits as i32, cap_height as i32, &mut h)
        });
        let mut cap = vec![<C::Hasher as Hasher<F>>::Hash::default(); 1 << cap_height];
        check(unsafe { sys::nlc_batch_cap(h, cap.as_mut_ptr() as *mut u64) });
        let handle = Arc::new(DeviceBatch(h));
        // the coefficients are the caller's own vectors; leaves / digests come back on first use of the field
        Self {
            polynomials,
            merkle_tree: MerkleTree::lazy(handle.clone(), MerkleCap(cap)),
            degree_log: log_n,
            rate_bits,
            blinding,
            handle,
        }
    }

    /// Fetches LDE values at the `index * step`th point.
    pub fn get_lde_values(&self, index: usize, step: usize) -> Vec<F> {
        let mut row = vec![F::ZERO; self.polynomials.len()];
        let idx = index as u64;
        check(unsafe { sys::nlc_batch_lde_rows(self.handle.0, &idx, 1, step, row.as_mut_ptr() as *mut u64) });
        row
    }

    /// Like `get_lde_values`, but fetches LDE values from a batch of `P::WIDTH` points, and returns packed values.
    pub fn get_lde_values_packed<P>(&self, index_start: usize, step: usize) -> Vec<P>
    where
        P: PackedField<Scalar = F>,
    {
        let w = self.polynomials.len();
        let mut rows = vec![F::ZERO; P::WIDTH * w];
        // one launch + one copy for the whole pack (a stripe of many packs: nlc_batch_lde_rows_range with count = k * WIDTH)
        check(unsafe { sys::nlc_batch_lde_rows_range(self.handle.0, index_start, P::WIDTH, step, rows.as_mut_ptr() as *mut u64) });
        (0..w).map(|j| *P::from_slice(&(0..P::WIDTH).map(|i| rows[i * w + j]).collect::<Vec<_>>())).collect()
    }

    /// Produces a batch opening proof.
    pub fn prove_openings(
        instance: &FriInstanceInfo<F, D>,
        oracles: &[&Self],
        challenger: &mut Challenger<F, C::Hasher>,
        fri_params: &FriParams,
        _timing: &mut TimingTree,
    ) -> FriProof<F, C::Hasher, D> {
        assert!(D == 2, "the device FRI works over the quadratic extension");
        // FriInstanceInfo -> nlc_fri_instance: points (D words each), batch offsets, (oracle, polynomial) per opening
        let mut points = Vec::new();
        let (mut offsets, mut poly_oracle, mut poly_index) = (vec![0u32], Vec::new(), Vec::new());
        for batch in &instance.batches {
            points.extend(batch.point.to_basefield_array().iter().map(|x| x.to_canonical_u64()));
            for info in &batch.polynomials {
                poly_oracle.push(info.oracle_index as u32);
                poly_index.push(info.polynomial_index as u32);
            }
            offsets.push(poly_oracle.len() as u32);
        }
        let ci = sys::nlc_fri_instance {
            n_batches: instance.batches.len() as i32,
            points: points.as_ptr(),
            batch_offsets: offsets.as_ptr(),
            poly_oracle: poly_oracle.as_ptr(),
            poly_index: poly_index.as_ptr(),
        };
        let mut p: sys::nlc_fri_params = unsafe { std::mem::zeroed() };
        p.degree_bits = fri_params.degree_bits as i32;
        p.rate_bits = fri_params.config.rate_bits as i32;
        p.cap_height = fri_params.config.cap_height as i32;
        p.proof_of_work_bits = fri_params.config.proof_of_work_bits as i32;
        p.num_query_rounds = fri_params.config.num_query_rounds as i32;
        p.n_reduction_rounds = fri_params.reduction_arity_bits.len() as i32;
        assert!(fri_params.reduction_arity_bits.len() <= sys::NLC_MAX_FRI_ROUNDS as usize);
        for (i, &b) in fri_params.reduction_arity_bits.iter().enumerate() {
            p.reduction_arity_bits[i] = b as i32;
        }
        let handles: Vec<*mut sys::nlc_batch> = oracles.iter().map(|o| o.handle.0).collect();
        // the Challenger's fields cross the ABI as data and come back advanced exactly as fri_proof advances them
        let mut ch = challenger.to_nlc();
        let words = unsafe { sys::nlc_fri_proof_words(&p, handles.as_ptr(), handles.len() as i32) };
        let mut flat = vec![0u64; words];
        check(unsafe { sys::nlc_prove_openings(ctx(), &ci, handles.as_ptr(), handles.len() as i32, &mut ch, &p, flat.as_mut_ptr()) });
        challenger.set_from_nlc(&ch);
        parse_fri_proof::<F, C, D>(&flat, fri_params, oracles)
    }
}

/// Slice the flat proof (layout documented at `nlc_fri_proof_words` in include/nlc_commit.h) into plonky2's FriProof.
fn parse_fri_proof<F: RichField + Extendable<D>, C: GenericConfig<D, F = F>, const D: usize>(
    f: &[u64],
    fp: &FriParams,
    oracles: &[&PolynomialBatch<F, C, D>],
) -> FriProof<F, C::Hasher, D> {
    let mut w = 0usize;
    let cap = fp.config.cap_height;
    let log_big_n = fp.degree_bits + fp.config.rate_bits;
    let hash = |f: &[u64], w: &mut usize| -> <C::Hasher as Hasher<F>>::Hash {
        let h = <C::Hasher as Hasher<F>>::Hash::from_u64s(&f[*w..*w + 4]);
        *w += 4;
        h
    };
    let ext = |f: &[u64], w: &mut usize| -> F::Extension {
        let e = F::Extension::from_basefield_array(std::array::from_fn(|k| F::from_canonical_u64(f[*w + k])));
        *w += D;
        e
    };
    let commit_phase_merkle_caps = (0..fp.reduction_arity_bits.len())
        .map(|_| MerkleCap((0..1 << cap).map(|_| hash(f, &mut w)).collect()))
        .collect();
    let final_poly = PolynomialCoeffs::new((0..fp.final_poly_len()).map(|_| ext(f, &mut w)).collect());
    let pow_witness = F::from_canonical_u64(f[w]);
    w += 1;
    let mut query_round_proofs = Vec::with_capacity(fp.config.num_query_rounds);
    for _ in 0..fp.config.num_query_rounds {
        let mut evals_proofs = Vec::with_capacity(oracles.len());
        for o in oracles {
            let n = o.polynomials.len();
            let leaf: Vec<F> = f[w..w + n].iter().map(|&x| F::from_canonical_u64(x)).collect();
            w += n;
            let siblings = (0..log_big_n - cap).map(|_| hash(f, &mut w)).collect();
            evals_proofs.push((leaf, MerkleProof { siblings }));
        }
        let mut lg = log_big_n;
        let mut steps = Vec::with_capacity(fp.reduction_arity_bits.len());
        for &ab in &fp.reduction_arity_bits {
            let evals = (0..1 << ab).map(|_| ext(f, &mut w)).collect();
            let siblings = (0..lg - ab - cap).map(|_| hash(f, &mut w)).collect();
            steps.push(FriQueryStep { evals, merkle_proof: MerkleProof { siblings } });
            lg -= ab;
        }
        query_round_proofs.push(FriQueryRound { initial_trees_proof: FriInitialTreeProof { evals_proofs }, steps });
    }
    assert_eq!(w, f.len(), "FriProof: flat layout mismatch");
    FriProof { commit_phase_merkle_caps, query_round_proofs, final_poly, pow_witness }
}

// Companion items of the patch (other files of the same crate):
//   hash/merkle_tree.rs   MerkleTree::{from_flat_parts, lazy}: `leaves` is kept as ONE flat N x W buffer (the library
//                         writes it in one piece) with `leaves: Vec<Vec<F>>` rebuilt from it only under the
//                         serializer; `lazy` fetches through nlc_batch_leaves / nlc_batch_digests on first access;
//                         MerkleTree::new(leaves, cap_height) -> nlc_merkle_new; MerkleTree::prove -> nlc_batch_open.
//   iop/challenger.rs     Challenger::{to_nlc, set_from_nlc}: sponge_state / input_buffer / output_buffer <-> nlc_challenger.
//   util/serialization    read_polynomial_batch -> nlc_batch_import(coeffs, .., expect_cap, expect_digests).
//   field/fft.rs, polynomial/mod.rs  fft / ifft / coset_fft / coset_ifft / lde_onto_coset above a size threshold ->
//                         nlc_fft / nlc_ifft / nlc_coset_fft / nlc_coset_ifft / nlc_lde_onto_coset.
