"""Command-line twins of the reference's operator scripts: same flags, same files.

    python -m volcanosv_b200.cli <script-name> <the reference script's argv>

e.g. `python -m volcanosv_b200.cli extract_reads_signature.py -i reads.bam -o out -chr 10`
replaces `python3 LI/extract_reads_signature.py -i reads.bam -o out -chr 10` (RVC:83-88).
`volcanosv-vc-large-indel.py` takes exactly the reference's flags (MAIN:10-36); `-cbam` is an optional
extension that hands over contig alignments made elsewhere instead of running RVC:49-58's command.
The executables under bin/VolcanoSV-vc/Large_INDEL/ carry the reference's script names and forward here.
"""
from __future__ import annotations

import argparse
import sys

from . import operators as OP


def _ers(argv):
    p = argparse.ArgumentParser(prog="extract_reads_signature.py")
    p.add_argument("--chr_number", "-chr", type=int)
    p.add_argument("--input_path", "-i")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--min_siglen", type=int, default=30)
    a, _ = p.parse_known_args(argv)
    OP.extract_reads_signature(a.input_path, a.output_dir, a.chr_number, a.min_siglen)


def _fpf(argv):
    p = argparse.ArgumentParser(prog="FP_filter_v1.py")
    p.add_argument("--input_path", "-i")
    p.add_argument("--signature_dir", "-sigd")
    p.add_argument("--output_path", "-o")
    a, _ = p.parse_known_args(argv)
    OP.fp_filter(a.input_path, a.signature_dir, a.output_path)


def _se(argv):
    p = argparse.ArgumentParser(prog="sig_extract.py")
    p.add_argument("input")
    p.add_argument("reference")
    p.add_argument("work_dir")
    p.add_argument("-b", "--batches", type=int, default=10_000_000)
    p.add_argument("-p", "--max_split_parts", type=int, default=7)
    p.add_argument("-q", "--min_mapq", type=int, default=20)
    p.add_argument("-r", "--min_read_len", type=int, default=500)
    p.add_argument("-include_bed", default=None)
    p.add_argument("-l", "--min_size", type=int, default=30)
    p.add_argument("-L", "--max_size", type=int, default=100000)
    p.add_argument("-sl", "--min_siglength", type=int, default=10)
    a, _ = p.parse_known_args(argv)          # cuteSV clustering flags are accepted and unused, as in the reference
    OP.sig_extract(a.input, a.reference, a.work_dir, a.include_bed, a.batches, a.min_mapq, a.min_read_len, a.min_siglength,
                   a.min_size, a.max_size, a.max_split_parts)


def _css(argv):
    p = argparse.ArgumentParser(prog="calculate_signature_support.py")
    p.add_argument("--vcffile", "-v")
    p.add_argument("--cutesv_dir", "-ct")
    p.add_argument("--flanking", "-f", type=int, default=1000)
    p.add_argument("--min_size", "-s", type=int, default=30)
    p.add_argument("--chr_num", "-chr", type=int, default=None)
    a, _ = p.parse_known_args(argv)
    OP.signature_support(a.vcffile, a.cutesv_dir, a.chr_num, a.flanking, a.min_size)


def _cov(argv):
    p = argparse.ArgumentParser(prog="filter_vcf_by_sig_cov_insdel.py")
    p.add_argument("--input_path", "-i")
    p.add_argument("--dtype", "-d")
    p.add_argument("--asm", "-a")
    p.add_argument("--vtype", "-v", choices=["INS", "DEL", "INSDEL"], default="INSDEL")
    a = p.parse_args(argv)
    OP.filter_by_sig_cov(a.input_path, a.dtype, a.asm, a.vtype)


def _gt(argv):
    p = argparse.ArgumentParser(prog="correct_gt_{del,ins}_real_data.py")
    p.add_argument("--input_path", "-i")
    p.add_argument("--output_path", "-o")
    p.add_argument("--bamfile", "-bam")
    p.add_argument("--sigfile", "--sig_file", "-sig")
    p.add_argument("--dtype", "-d", choices=["ONT", "Hifi", "CLR"])
    p.add_argument("--vtype", "-v", choices=["INS", "DEL"])
    a, _ = p.parse_known_args(argv)
    OP.correct_gt(a.input_path, a.output_path, a.bamfile, a.sigfile, a.dtype, a.vtype)


def _rr(argv):
    p = argparse.ArgumentParser(prog="remove_redundancy.py")
    p.add_argument("--input_path", "-i")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--dist_thresh", "-r", type=int, default=500)
    p.add_argument("--dist_thresh_del", "-rd", type=int, default=3000)
    p.add_argument("--overlap_thresh", "-O", type=float, default=0)
    p.add_argument("--size_sim_thresh", "-P", type=float, default=0.5)
    p.add_argument("--size_sim_thresh_del", "-Pd", type=float, default=0.1)
    p.add_argument("--seq_sim_thresh", "-p", type=float, default=0.5)
    a, _ = p.parse_known_args(argv)
    OP.remove_redundancy(a.input_path, a.output_dir, a.dist_thresh, a.dist_thresh_del, a.overlap_thresh, a.size_sim_thresh,
                         a.size_sim_thresh_del, a.seq_sim_thresh)


def _ecs(dtype):
    def run(argv):
        p = argparse.ArgumentParser(prog="extract_contig_signature_%s.py" % dtype)
        p.add_argument("--chr_number", "-chr", type=int)
        p.add_argument("--bam_path", "-bam")
        p.add_argument("--contig_path", "-contig")
        p.add_argument("--header_path", "-header")
        p.add_argument("--ref_path", "-ref")
        p.add_argument("--output_dir", "-o")
        p.add_argument("--min_cigar_mapq", type=int, default=50)
        p.add_argument("--min_split_mapq", type=int, default=50)
        a, _ = p.parse_known_args(argv)
        # the reference assigns min_split_mapq = args.min_cigar_mapq (ECS_H:41)
        OP.extract_contig_signature(a.bam_path, a.contig_path, a.ref_path, a.output_dir, a.chr_number, dtype, a.header_path,
                                    a.min_cigar_mapq, a.min_cigar_mapq)
    return run


def _fgc(argv):
    p = argparse.ArgumentParser(prog="filter_GT_correction.py")
    p.add_argument("--vcffile", "-vcf")
    p.add_argument("--bamfile", "-bam")
    p.add_argument("--reference", "-ref")
    p.add_argument("--pre_cutesig", "-presig")
    p.add_argument("--dtype", "-dtype", choices=["Hifi", "CLR", "ONT"])
    p.add_argument("--chr_num", "-chr", type=int, default=None)
    p.add_argument("--n_thread", "-t", type=int, default=22)
    a = p.parse_args(argv)
    OP.filter_gt_correction(a.vcffile, a.bamfile, a.reference, a.dtype, a.chr_num, a.pre_cutesig)


def _main(argv):
    p = argparse.ArgumentParser(prog="volcanosv-vc-large-indel.py")
    p.add_argument("--input_dir", "-i")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--data_type", "-dtype")
    p.add_argument("--bam_file", "-bam")
    p.add_argument("--reference", "-ref")
    p.add_argument("--read_signature_dir", "-rdsig")
    p.add_argument("--pre_cutesig", "-presig")
    p.add_argument("--chr_num", "-chr", type=int, default=None)
    p.add_argument("--n_thread", "-t", type=int, default=11)
    p.add_argument("--n_thread_align", "-ta", type=int, default=10)
    p.add_argument("--mem_per_thread", "-mempt", default="768M")
    p.add_argument("--prefix", "-px", default="Sample")
    p.add_argument("--contig_bam", "-cbam", default=None,
                   help="OPTIONAL extension: sorted contig-to-reference alignments made elsewhere (whole-genome mode: a pattern with "
                        "{chr}).  Without it the contigs are aligned exactly as Raw_variant_call.py:49-58 does (minimap2 | samtools sort)")
    a = p.parse_args(argv)
    if a.chr_num is None:                      # whole-genome mode (MAIN:266-279); one process per GPU under torchrun
        rank, world, barrier = _ranks()
        out = OP.large_indel_wgs(a.contig_bam, a.input_dir, a.reference, a.bam_file, a.output_dir, a.data_type, a.prefix,
                                 a.read_signature_dir, a.pre_cutesig, rank=rank, world=world, barrier=barrier,
                                 n_thread_align=a.n_thread_align)
        if out:
            print(out)
        return
    contigs = "%s/chr%d/assembly/final_contigs/%s_final_contigs.fa" % (a.input_dir, a.chr_num, a.prefix)
    print(OP.large_indel_chr(a.contig_bam, contigs, a.reference, a.bam_file, a.output_dir, a.data_type, a.chr_num, a.prefix,
                             a.read_signature_dir, a.pre_cutesig, n_thread_align=a.n_thread_align))


def _rvc(argv):
    p = argparse.ArgumentParser(prog="Raw_variant_call.py")
    p.add_argument("--contig_path", "-contig")
    p.add_argument("--reference_path", "-ref")
    p.add_argument("--signature_dir", "-sigd")
    p.add_argument("--rbam_file", "-rbam")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--data_type", "-dtype")
    p.add_argument("--chr_num", "-chr", type=int)
    p.add_argument("--header_file", "-header", default=None)
    p.add_argument("--n_thread", "-t", type=int, default=10)
    p.add_argument("--mem_per_thread", "-mempt", default="1G")
    a, _ = p.parse_known_args(argv)
    print(OP.raw_variant_call(None, a.contig_path, a.reference_path, a.output_dir, a.data_type, a.chr_num, a.rbam_file,
                              a.signature_dir, a.header_file, n_thread=a.n_thread, mem_per_thread=a.mem_per_thread))


def _rvc_otherasm(argv):
    p = argparse.ArgumentParser(prog="Raw_variant_call-otherasm.py")
    p.add_argument("--contig_path", "-contig")
    p.add_argument("--contig_bam", "-cbam")
    p.add_argument("--reference_path", "-ref")
    p.add_argument("--rbam_file", "-rbam")
    p.add_argument("--signature_dir", "-sigd")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--data_type", "-dtype")
    p.add_argument("--n_thread", "-t", type=int, default=10)
    p.add_argument("--chr_num", "-chr", type=int)
    p.add_argument("--header_file", "-header", default=None)
    a, _ = p.parse_known_args(argv)
    print(OP.raw_variant_call(a.contig_bam, a.contig_path, a.reference_path, a.output_dir, a.data_type, a.chr_num, a.rbam_file,
                              a.signature_dir, a.header_file))


def _ranks():
    import os
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    barrier = None
    if world > 1:
        import torch
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        from .shard import bind_host_to_gpu
        bind_host_to_gpu(local)            # staging buffers and feeder threads on the GPU's own socket
        dist.init_process_group("gloo")    # file-level hand-over only: no data-path collective
        barrier = dist.barrier
    return rank, world, barrier


def _main_otherasm(argv):
    p = argparse.ArgumentParser(prog="volcanosv-vc-large-indel-otherasm.py")
    p.add_argument("--input_path", "-i")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--data_type", "-dtype")
    p.add_argument("--bam_file", "-bam")
    p.add_argument("--reference", "-ref")
    p.add_argument("--read_signature_dir", "-rdsig")
    p.add_argument("--pre_cutesig", "-presig")
    p.add_argument("--n_thread", "-t", type=int, default=11)
    p.add_argument("--prefix", "-px", default="Sample")
    p.add_argument("--n_thread_align", "-ta", type=int, default=10)
    p.add_argument("--mem_per_thread", "-mempt", default="768M")
    p.add_argument("--contig_bam", "-cbam", default=None,
                   help="OPTIONAL extension: sorted whole-genome contig alignments made elsewhere; without it otherasm:232-245's "
                        "command runs (minimap2 -x asm20 | samtools sort)")
    a, _ = p.parse_known_args(argv)
    rank, world, barrier = _ranks()
    out = OP.large_indel_otherasm(a.input_path, a.contig_bam, a.reference, a.bam_file, a.output_dir, a.data_type, a.prefix,
                                  a.read_signature_dir, a.pre_cutesig, rank=rank, world=world, barrier=barrier,
                                  n_thread=a.n_thread, mem_per_thread=a.mem_per_thread)
    if out:
        print(out)


def _main_otherasm_2hap(argv):
    p = argparse.ArgumentParser(prog="volcanosv-vc-large-indel-otherasm-2hap.py")
    p.add_argument("--hp1_fa", "-hp1")
    p.add_argument("--hp2_fa", "-hp2")
    p.add_argument("--output_dir", "-o")
    p.add_argument("--data_type", "-dtype")
    p.add_argument("--bam_file", "-bam")
    p.add_argument("--reference", "-ref")
    p.add_argument("--read_signature_dir", "-rdsig")
    p.add_argument("--pre_cutesig", "-presig")
    p.add_argument("--n_thread", "-t", type=int, default=11)
    p.add_argument("--prefix", "-px", default="Sample")
    p.add_argument("--n_thread_align", "-ta", type=int, default=10)
    p.add_argument("--mem_per_thread", "-mempt", default="768M")
    p.add_argument("--contig_bam", "-cbam", default=None,
                   help="OPTIONAL extension: sorted merge of both haplotypes' contig alignments made elsewhere; without it 2hap:250-258's "
                        "commands run")
    a, _ = p.parse_known_args(argv)
    rank, world, barrier = _ranks()
    out = OP.large_indel_otherasm_2hap(a.hp1_fa, a.hp2_fa, a.contig_bam, a.reference, a.bam_file, a.output_dir, a.data_type,
                                       a.prefix, a.read_signature_dir, a.pre_cutesig, rank=rank, world=world, barrier=barrier,
                                       n_thread=a.n_thread, mem_per_thread=a.mem_per_thread)
    if out:
        print(out)


COMMANDS = {
    "extract_reads_signature.py": _ers, "FP_filter_v1.py": _fpf, "sig_extract.py": _se,
    "calculate_signature_support.py": _css, "filter_vcf_by_sig_cov_insdel.py": _cov,
    "correct_gt_del_real_data.py": _gt, "correct_gt_ins_real_data.py": _gt, "remove_redundancy.py": _rr,
    "extract_contig_signature_Hifi.py": _ecs("Hifi"), "extract_contig_signature_ONT.py": _ecs("ONT"),
    "extract_contig_signature_CLR.py": _ecs("CLR"), "filter_GT_correction.py": _fgc,
    "volcanosv-vc-large-indel.py": _main,
    "Raw_variant_call.py": _rvc,
    "Raw_variant_call-otherasm.py": _rvc_otherasm,
    "volcanosv-vc-large-indel-otherasm.py": _main_otherasm,
    "volcanosv-vc-large-indel-otherasm-2hap.py": _main_otherasm_2hap,
}


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in COMMANDS:
        sys.exit("usage: python -m volcanosv_b200.cli <%s> [the script's arguments]" % " | ".join(COMMANDS))
    COMMANDS[argv[0]](argv[1:])


if __name__ == "__main__":
    main()
