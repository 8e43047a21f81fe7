#!/usr/bin/perl
# TEST INFRASTRUCTURE ONLY.  Runs the UNMODIFIED scoring loop of the reference's
# perl/run_rankflankmap.pl (the `foreach my $file (@allfiles) { ... }` block, lines 97-231 at the
# surveyed commit) on a folder of .map files, without a database: the loop's text is read from the
# reference script where it lies (nothing is copied into this repository) and evaluated here with
# the two database helpers it calls replaced by stubs and the batch size set so high that the loop
# never flushes -- the rows it would have inserted are then still in @map_rows / @rankflank_rows
# and are printed as the tab-separated files the library writes.
#
# usage: rankflank_harness.pl REFERENCE_SCRIPT MAPDIR RANK_TSV MAP_TSV
# The reference takes the files in readdir order; so that outputs can be compared byte for byte the
# harness hands it the names sorted (the library's documented order).
use strict;
use warnings;

my ( $script, $mapdir_arg, $rank_tsv, $map_tsv ) = @ARGV;
die "usage: rankflank_harness.pl REFERENCE_SCRIPT MAPDIR RANK_TSV MAP_TSV\n" unless defined $map_tsv;
open( my $sf, '<', $script ) or die "cannot read $script: $!";
my $src = do { local $/; <$sf> };
close($sf);
$src =~ /\n(foreach my \$file \(\@allfiles\) \{.*?\n\}[^\n]*# end of foreach \@files)/s
    or die "the scoring loop was not found in $script\n";
my $loop = $1;

# what the loop expects from the part of the script before it (run_rankflankmap.pl:40-95)
our $mapdir = $mapdir_arg;
opendir( my $dh, $mapdir ) or die $!;
our @allfiles = sort grep { -f "$mapdir/$_" } readdir($dh);
closedir($dh);
our ( $clusters_processed, $j, $k, $uploadedrank, $uploadedmap ) = ( 0, 0, 0, 0, 0 );
our ( @map_rows, @rankflank_rows );
our ( $dbh, $map_insert_sth, $rankflank_insert_sth );
our $RECORDS_PER_INFILE_INSERT = 1e15;    # never reached: nothing is flushed inside the loop
sub gen_exec_array_cb { die "flush inside the loop\n" }
sub vs_db_insert      { die "flush inside the loop\n" }

# the reference declares these with `my` at file scope; here they are package variables
$loop =~ s/^/no strict 'vars'; no warnings; /;
eval $loop;
die "reference loop failed: $@" if $@;

open( my $fr, '>', $rank_tsv ) or die $!;
open( my $fm, '>', $map_tsv )  or die $!;
print $fm join( "\t", @$_ ), "\n" for @map_rows;
print $fr join( "\t", @$_ ), "\n" for @rankflank_rows;
close($fr);
close($fm);
